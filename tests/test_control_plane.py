"""Control-plane parity with the reference (goldens from SURVEY.md Appendix A, plus a live
differential test against /root/reference when it is mounted)."""
import importlib.util
import json
import os
import statistics
import sys
from pathlib import Path

import pytest
from hypothesis import given, settings, strategies as st

from distributed_llm_training_gpu_manager_b200.ai_engine import (DeepSpeedConfig, DeepSpeedLauncher, GPUManager,
                                                                  HaltAction, HaltPolicy, LossSpikeMonitor,
                                                                  MonitorConfig, OffloadDevice, TrainingMetrics)
from distributed_llm_training_gpu_manager_b200.ai_engine.loss_monitor import RollingStats

REF = Path("/root/reference/ai_engine")


def _ref(mod):
    if not (REF / f"{mod}.py").exists():
        pytest.skip("reference not mounted")
    spec = importlib.util.spec_from_file_location(f"ref_{mod}", REF / f"{mod}.py")
    m = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = m
    spec.loader.exec_module(m)
    return m


# ----------------------------------------------------------------------------- launcher
def test_default_config_golden():
    cfg = DeepSpeedLauncher().generate_config(DeepSpeedConfig())
    assert list(cfg) == ["train_micro_batch_size_per_gpu", "gradient_accumulation_steps", "gradient_clipping",
                         "steps_per_print", "wall_clock_breakdown", "dump_state", "zero_optimization", "scheduler",
                         "optimizer", "communication_data_type", "prescale_gradients", "gradient_predivide_factor",
                         "fp16", "activation_checkpointing"]
    assert cfg["zero_optimization"] == {
        "stage": 3, "allgather_partitions": True, "allgather_bucket_size": 500000000, "overlap_comm": True,
        "reduce_scatter": True, "reduce_bucket_size": 500000000, "contiguous_gradients": True,
        "round_robin_gradients": True, "stage3_max_live_parameters": 1000000000,
        "stage3_max_reuse_distance": 1000000000, "stage3_prefetch_bucket_size": 500000000,
        "stage3_param_persistence_threshold": 100000, "stage3_gather_16bit_weights_on_model_save": True,
        "sub_group_size": 1000000000,
        "offload_optimizer": {"device": "cpu", "pin_memory": True, "buffer_count": 4, "fast_init": False},
        "offload_param": {"device": "cpu", "pin_memory": True, "buffer_count": 5, "buffer_size": 1000000000,
                          "max_in_cpu": 1000000000}}
    assert cfg["fp16"] == {"enabled": True, "loss_scale": 0, "loss_scale_window": 1000, "initial_scale_power": 16,
                           "hysteresis": 2, "min_loss_scale": 1}
    assert cfg["steps_per_print"] == 100 and cfg["communication_data_type"] == "fp16"


@pytest.mark.parametrize("stage", [0, 1, 2, 3])
@pytest.mark.parametrize("bf16,fp16", [(True, True), (False, True), (False, False)])
@pytest.mark.parametrize("off", ["none", "cpu", "nvme"])
@pytest.mark.parametrize("elastic", [False, True])
def test_config_matches_reference(stage, bf16, fp16, off, elastic):
    ref = _ref("deepspeed_launcher")
    kw = dict(zero_stage=stage, bf16_enabled=bf16, fp16_enabled=fp16, offload_optimizer=off, offload_param=off,
              elastic_training=elastic, num_gpus=8, num_nodes=2, activation_checkpointing=(stage != 1))
    mine = DeepSpeedLauncher().generate_config(DeepSpeedConfig(**kw))
    theirs = ref.DeepSpeedLauncher().generate_config(ref.DeepSpeedConfig(**kw))
    assert json.dumps(mine) == json.dumps(theirs)  # same keys, same order, same values


def test_elastic_block_and_precedence():
    cfg = DeepSpeedLauncher().generate_config(DeepSpeedConfig(elastic_training=True, num_gpus=8, num_nodes=2,
                                                              bf16_enabled=True))
    assert cfg["elasticity"] == {"enabled": True, "max_train_batch_size": 512, "min_train_batch_size": 4,
                                 "min_gpus": 1, "max_gpus": 16, "prefer_larger_batch_size": True}
    assert "bf16" in cfg and "fp16" not in cfg
    c2 = DeepSpeedLauncher().generate_config(DeepSpeedConfig(zero_stage=1))
    assert "stage3_max_live_parameters" not in c2["zero_optimization"] and "offload_param" in c2["zero_optimization"]


def test_commands_and_dry_run(tmp_path):
    ds = DeepSpeedLauncher(engine="deepspeed")
    one = ds.build_launch_command(DeepSpeedConfig(num_gpus=8), "train.py", config_path="/tmp/x.json")
    assert one == "deepspeed --num_gpus=8 train.py --deepspeed_config=/tmp/x.json"
    multi = ds.build_launch_command(DeepSpeedConfig(num_gpus=8, num_nodes=2, master_addr="10.0.0.1"), "train.py",
                                    ["--a", "b", "c"], "/tmp/x.json")
    assert multi == ("deepspeed --num_nodes=2 --master_addr=10.0.0.1 --master_port=29500 --num_gpus=8 train.py "
                     "--deepspeed_config=/tmp/x.json --a b c")
    native = DeepSpeedLauncher()
    argv = native.build_launch_argv(DeepSpeedConfig(num_gpus=2), "train.py", ["--name", "has space"], "/tmp/x.json")
    assert argv[1:3] == ["-m", "distributed_llm_training_gpu_manager_b200.engine.launch"]
    assert argv[-1] == "has space"                       # not mangled (reference splits on whitespace)
    res = native.launch(DeepSpeedConfig(num_gpus=8, num_nodes=2), "train.py", dry_run=True)
    assert res.status == "dry_run" and res.num_gpus == 16 and res.effective_batch_size == 4 * 8 * 8 * 2
    assert res.details["precision"] == "fp16" and os.path.exists(res.config_path)
    a = native.write_config(DeepSpeedConfig(model_name="same"))
    b = native.write_config(DeepSpeedConfig(model_name="same"))
    assert a != b                                        # same-second collision does not overwrite
    failed = DeepSpeedLauncher(deepspeed_path="/nonexistent/deepspeed").launch(DeepSpeedConfig(), "t.py")
    assert failed.status == "failed" and "No such file" in failed.details["error"]


def test_presets():
    p = DeepSpeedLauncher.presets()
    assert set(p) == {"7b", "13b", "70b"}
    assert (p["7b"].train_micro_batch_size_per_gpu, p["7b"].gradient_accumulation_steps, p["7b"].num_gpus) == (2, 16, 4)
    assert p["7b"].offload_param is OffloadDevice.NONE and p["13b"].offload_param is OffloadDevice.CPU
    assert p["70b"].bf16_enabled and not p["70b"].fp16_enabled and p["70b"].cpu_checkpointing and p["70b"].num_nodes == 2
    assert p["70b"].effective_batch_size == 1024 and p["13b"].effective_batch_size == 256
    assert "llama3-8b" in DeepSpeedLauncher.b200_presets()


def test_job_registry_runs_and_stops(tmp_path):
    from distributed_llm_training_gpu_manager_b200.ai_engine import JobRegistry
    reg = JobRegistry(str(tmp_path))
    rec = reg.spawn("j1", [sys.executable, "-c", "print('hello'); import time; time.sleep(60)"], dict(os.environ))
    assert reg.get("j1").status == "running"
    import time; time.sleep(0.5)
    out = reg.stop("j1", grace_s=5)
    assert out.status == "stopped" and out.returncode is not None
    assert "hello" in reg.tail("j1")
    rec2 = reg.spawn("j2", [sys.executable, "-c", "raise SystemExit(3)"], dict(os.environ))
    assert reg.wait("j2", 20).status == "failed" and reg.get("j2").returncode == 3


# ----------------------------------------------------------------------------- loss monitor
def _feed(mon, losses, start=0, **kw):
    out = []
    for i, l in enumerate(losses):
        out.append(mon.ingest(TrainingMetrics(step=start + i, loss=l, **kw)))
    return out


def test_monitor_appendix_a3_semantics():
    m = LossSpikeMonitor()
    assert all(not a for a in _feed(m, [2.5] * 9 + [100.0]))           # <10 samples: no spike check
    m = LossSpikeMonitor()
    _feed(m, [2.5] * 10)
    a = m.ingest(TrainingMetrics(step=10, loss=2.5 + 1e-6))              # sigma=0 floor 1e-8 -> CRITICAL
    assert [x.alert_type for x in a] == ["spike"] and a[0].severity.value == "critical"
    assert not m.ingest(TrainingMetrics(step=11, loss=50.0))             # cooldown 20 steps
    assert [x.alert_type for x in m.ingest(TrainingMetrics(step=30, loss=500.0))] == ["spike"]
    # NaN: returned, not stored, window untouched, steps counted
    m = LossSpikeMonitor()
    _feed(m, [1.0] * 3)
    a = m.ingest(TrainingMetrics(step=3, loss=float("nan")))
    assert a[0].alert_type == "divergence" and m.state.alert_count == 0 and m.state.total_steps == 4
    assert len(m._loss_history) == 3
    # divergence has no cooldown; >1e6 can fire divergence and spike together
    m = LossSpikeMonitor()
    _feed(m, [1.0] * 10)
    a = m.ingest(TrainingMetrics(step=10, loss=2e6))
    assert sorted(x.alert_type for x in a) == ["divergence", "spike"]
    assert [x.alert_type for x in m.ingest(TrainingMetrics(step=11, loss=3e6))] == ["divergence"]
    # plateau cadence with patience 5, cooldown 2
    m = LossSpikeMonitor(MonitorConfig(plateau_patience=5, cooldown_steps=2))
    fired = [i for i, a in enumerate(_feed(m, [1.0] * 13)) if any(x.alert_type == "plateau" for x in a)]
    assert fired == [5, 7, 9, 11]
    # NaN grad norm ignored, Inf triggers
    m = LossSpikeMonitor()
    assert not m.ingest(TrainingMetrics(step=0, loss=1.0, gradient_norm=float("nan")))
    assert m.ingest(TrainingMetrics(step=1, loss=0.9, gradient_norm=float("inf")))[0].alert_type == "gradient_explosion"
    # fresh summary is JSON-safe through the API layer, raw best_loss is inf
    assert LossSpikeMonitor().get_summary()["best_loss"] == float("inf")


@settings(max_examples=60, deadline=None)
@given(st.lists(st.floats(min_value=-1e3, max_value=1e3, allow_nan=False), min_size=1, max_size=400),
       st.integers(min_value=2, max_value=50))
def test_rolling_stats_match_statistics(values, window):
    r = RollingStats(window)
    for i, v in enumerate(values):
        r.push(v)
        tail = values[max(0, i + 1 - window): i + 1]
        assert r.mean == pytest.approx(statistics.mean(tail), rel=1e-9, abs=1e-9)
        if len(tail) > 1:
            assert r.std == pytest.approx(statistics.stdev(tail), rel=1e-6, abs=1e-7)


@settings(max_examples=25, deadline=None)
@given(st.lists(st.tuples(st.floats(min_value=0.0, max_value=20.0), st.floats(min_value=0.0, max_value=300.0),
                          st.floats(min_value=0.0, max_value=1e-2)), min_size=1, max_size=150))
def test_monitor_matches_reference_alert_stream(seq):
    ref = _ref("loss_monitor")
    cfg = dict(window_size=20, plateau_patience=15, cooldown_steps=5)
    mine, theirs = LossSpikeMonitor(MonitorConfig(**cfg)), ref.LossSpikeMonitor(ref.MonitorConfig(**cfg))
    for i, (l, g, lr) in enumerate(seq):
        a = mine.ingest(TrainingMetrics(step=i, loss=l, gradient_norm=g, learning_rate=lr))
        b = theirs.ingest(ref.TrainingMetrics(step=i, loss=l, gradient_norm=g, learning_rate=lr))
        assert [(x.alert_type, x.severity.value) for x in a] == [(x.alert_type, x.severity.value) for x in b]
    assert mine.state.alert_count == theirs.state.alert_count


def test_monitor_additions():
    m = LossSpikeMonitor(MonitorConfig(max_alerts_per_type=2, cooldown_steps=0, window_size=10))
    _feed(m, [1.0] * 10)
    for i in range(5):
        m.ingest(TrainingMetrics(step=10 + i, loss=1e9))
    assert m.get_summary()["alert_counts_by_type"]["divergence"] == 2      # cap honoured
    sd = m.state_dict()
    m2 = LossSpikeMonitor(MonitorConfig(window_size=10)); m2.load_state_dict(json.loads(json.dumps(sd)))
    assert list(m2._loss_history) == list(m._loss_history) and m2.state.total_steps == m.state.total_steps
    a = m2.ingest_device_flags(step=99, loss=1.0, nonfinite_grads=True, grad_norm_sq=float("nan"))
    assert a[0].alert_type == "divergence"
    assert HaltPolicy().decide(a) is HaltAction.ROLLBACK
    assert HaltPolicy().decide([]) is HaltAction.CONTINUE


# ----------------------------------------------------------------------------- gpu manager
B200_XML = """<?xml version="1.0" ?><nvidia_smi_log><driver_version>580.159</driver_version><cuda_version>12.9</cuda_version>
<gpu id="00000000:1B:00.0"><product_name>NVIDIA B200</product_name><uuid>GPU-aaa</uuid><pci><pci_bus_id>00000000:1B:00.0</pci_bus_id></pci>
<fan_speed>N/A</fan_speed><fb_memory_usage><total>183359 MiB</total><used>175000 MiB</used><free>8359 MiB</free></fb_memory_usage>
<compute_mode>Default</compute_mode><utilization><gpu_util>97 %</gpu_util><memory_util>60 %</memory_util></utilization>
<temperature><gpu_temp>91 C</gpu_temp></temperature>
<gpu_power_readings><power_draw>950.00 W</power_draw><current_power_limit>1000.00 W</current_power_limit></gpu_power_readings>
<processes><process_info><pid>4242</pid><process_name>python</process_name><used_memory>170000 MiB</used_memory></process_info></processes></gpu>
<gpu id="00000000:2B:00.0"><product_name>NVIDIA B200</product_name><uuid>GPU-bbb</uuid>
<fb_memory_usage><total>183359 MiB</total><used>1000 MiB</used><free>182359 MiB</free></fb_memory_usage>
<utilization><gpu_util>3 %</gpu_util></utilization><temperature><gpu_temp>40 C</gpu_temp></temperature>
<gpu_power_readings><power_draw>140.00 W</power_draw><current_power_limit>1000.00 W</current_power_limit></gpu_power_readings></gpu>
</nvidia_smi_log>"""


def test_gpu_manager_xml_csv_and_selection():
    gm = GPUManager(backend="smi")
    d0, d1 = gm.parse_xml(B200_XML)
    assert d0.health.value == "critical" and d0.alerts[:3] == [
        "CRITICAL: Temperature 91°C exceeds 90°C", "CRITICAL: Memory 95% exceeds 95%",
        "WARNING: Utilization 97% at max capacity"]
    assert d0.power_limit_watts == 1000.0 and "WARNING: Power 950W near limit 1000W" in d0.alerts  # fixed tag
    assert d0.fan_speed_pct == 0 and d0.processes[0].pid == 4242 and d0.driver_version == "580.159"
    assert d1.health.value == "healthy" and d1.is_available and not d0.is_available
    fleet = gm.get_fleet_status(B200_XML)
    assert fleet.total_gpus == 2 and fleet.available_gpus == 1 and fleet.source == "xml"
    assert fleet.alerts[0].startswith("GPU 0 (NVIDIA B200): CRITICAL")
    assert gm.select_best_gpu(1000, B200_XML).index == 1 and gm.select_best_gpu(10**7, B200_XML) is None
    csv = ("0, NVIDIA B200, GPU-aaa, 91, 97, 60, 175000, 183359, 8359, 950.0, 1000.0, [N/A]\n"
           "1, NVIDIA B200, GPU-bbb, 40, 3, 1, 1000, 183359, 182359, 140.0, 1000.0, [N/A]\nshort,line\n")
    c0, c1 = gm.parse_csv(csv)
    assert "WARNING: Power 950W near limit 1000W" in c0.alerts and c1.memory_free_mib == 182359
    empty = GPUManager(nvidia_smi_path="/nonexistent/nvidia-smi", backend="smi").get_fleet_status()
    assert empty.total_gpus == 0 and empty.alerts == ["Unable to query nvidia-smi. No GPUs detected."]
    with pytest.raises(RuntimeError):
        GPUManager(nvidia_smi_path="/nonexistent/nvidia-smi", backend="smi").select_best_gpu()
    mock = gm.get_mock_fleet()
    assert mock.total_gpus == 2 and mock.devices[1].alerts == ["WARNING: Memory 86.7% exceeds 85%"]


def test_multi_gpu_placement_prefers_connected_clique():
    gm = GPUManager(backend="smi")
    devs = gm.parse_xml(B200_XML.replace("175000 MiB", "1000 MiB").replace("91 C", "50 C").replace("97 %", "5 %"))
    assert [d.index for d in gm.select_gpus(2, devices=devs)] == [0, 1]
    assert gm.select_gpus(3, devices=devs) is None
    from distributed_llm_training_gpu_manager_b200.ai_engine.gpu_manager import GPUDevice
    four = [GPUDevice(index=i, memory_total_mib=100, memory_free_mib=100 - i) for i in range(4)]
    links = {0: {1: 0, 2: 4, 3: 0}, 1: {0: 0, 2: 0, 3: 4}, 2: {0: 4, 1: 0, 3: 0}, 3: {0: 0, 1: 4, 2: 0}}
    assert [d.index for d in gm.select_gpus(2, devices=four, link_matrix=links)] == [0, 2]
    four[0].throttle_reasons = ["hw_thermal_slowdown"]
    assert [d.index for d in gm.select_gpus(2, devices=four, link_matrix=links)] == [1, 3]


def test_topology_mock_uniform_fabric():
    from distributed_llm_training_gpu_manager_b200.parallel.topology import mock_topology
    t = mock_topology()
    assert t.node_type == "HGX B200 8-GPU" and t.interconnect == "NVSwitch Fabric"
    assert t.matrix_labels()["GPU_0"]["GPU_7"] == "NV18" and t.bottlenecks() == [] and t.rank_order() == list(range(8))


# ----------------------------------------------------------------------------- spot resiliency
def test_spot_manager_detects_and_checkpoints(tmp_path):
    import asyncio, http.server, threading
    from distributed_llm_training_gpu_manager_b200.ai_engine import JobRegistry, SpotInstanceResiliencyManager
    from distributed_llm_training_gpu_manager_b200.ai_engine.spot_resiliency import AwsSpotProbe, GcpPreemptProbe

    state = {"armed": False}

    class H(http.server.BaseHTTPRequestHandler):
        def log_message(self, *a): pass
        def do_PUT(self):
            self.send_response(200); self.end_headers(); self.wfile.write(b"tok")
        def do_GET(self):
            if self.path.endswith("instance-action") and state["armed"]:
                self.send_response(200); self.end_headers(); self.wfile.write(b'{"action":"terminate"}')
            elif self.path.endswith("preempted"):
                self.send_response(200); self.end_headers(); self.wfile.write(b"TRUE" if state["armed"] else b"FALSE")
            else:
                self.send_response(404); self.end_headers()

    srv = http.server.HTTPServer(("127.0.0.1", 0), H)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    base = f"http://127.0.0.1:{srv.server_port}"
    aws = AwsSpotProbe(f"{base}/latest/meta-data/spot/instance-action", f"{base}/latest/api/token")
    gcp = GcpPreemptProbe(f"{base}/computeMetadata/v1/instance/preempted")
    assert aws() is None and gcp() is None
    reg = JobRegistry(str(tmp_path))
    script = ("import signal,sys,time\n"
              "signal.signal(signal.SIGTERM, lambda *a: (open(sys.argv[1],'w').write('ckpt'), sys.exit(0)))\n"
              "print('up', flush=True)\nwhile True: time.sleep(0.05)\n")
    marker = tmp_path / "emergency.ckpt"
    reg.spawn("job", [sys.executable, "-c", script, str(marker)], dict(os.environ))
    import time; time.sleep(0.7)

    class L:  # launcher-shaped object
        registry = reg

    mgr = SpotInstanceResiliencyManager(check_interval_sec=0.05, probes=[aws, gcp], grace_seconds=10)

    async def go():
        task = asyncio.create_task(mgr.monitor_preemption_notices(L()))
        await asyncio.sleep(0.3)
        assert not task.done()                       # nothing fires until a notice exists
        state["armed"] = True
        return await asyncio.wait_for(task, 20)

    out = asyncio.run(go())
    srv.shutdown()
    assert out["notice"] == "aws" and out["jobs"]["job"]["checkpointed"] is True
    assert marker.read_text() == "ckpt" and mgr.last_notice.source == "aws"


# ----------------------------------------------------------------------------- HTTP API
def test_http_surface():
    from fastapi.testclient import TestClient
    from distributed_llm_training_gpu_manager_b200.backend.main import app
    c = TestClient(app)
    assert c.get("/").json()["name"] == "MLOps Platform API" and c.get("/health").json() == {"status": "healthy"}
    paths = {r.path for r in app.routes}
    for p in ["/api/v1/gpu/fleet", "/api/v1/gpu/fleet/mock", "/api/v1/gpu/select", "/api/v1/gpu/devices/{index}",
              "/api/v1/gpu/alerts", "/api/v1/training/launch", "/api/v1/training/launch/preset",
              "/api/v1/training/presets", "/api/v1/training/config/generate", "/api/v1/monitoring/create",
              "/api/v1/monitoring/ingest", "/api/v1/monitoring/ingest/single", "/api/v1/monitoring/summary/{job_id}",
              "/api/v1/monitoring/loss-curve/{job_id}", "/api/v1/monitoring/reset/{job_id}", "/api/v1/monitoring/jobs",
              "/api/v1/nvlink/topology", "/api/v1/jobs/{job_id}", "/metrics"]:
        assert p in paths, p
    assert c.get("/api/v1/gpu/fleet/mock").json()["total_gpus"] == 2
    assert c.get("/api/v1/gpu/devices/99").status_code == 404
    r = c.post("/api/v1/training/launch", json={"training_script": "t.py", "num_gpus": 8, "num_nodes": 2})
    assert r.status_code == 200 and r.json()["status"] == "dry_run" and r.json()["effective_batch_size"] == 512
    assert c.post("/api/v1/training/launch", json={"training_script": "t.py", "offload_optimizer": "gpu"}).status_code == 422
    assert c.post("/api/v1/training/launch/preset", json={"preset": "nope", "training_script": "t.py"}).status_code == 400
    assert set(c.get("/api/v1/training/presets").json()) >= {"7b", "13b", "70b"}
    g = c.post("/api/v1/training/config/generate", json={"training_script": "t.py", "zero_stage": 2, "bf16": True}).json()
    assert g["deepspeed_config"]["zero_optimization"]["stage"] == 2 and g["effective_batch_size"] == 32
    assert "native_engine_handling" not in g                      # same payload as the reference unless asked
    ge = c.post("/api/v1/training/config/generate?explain=true", json={"training_script": "t.py", "zero_stage": 3}).json()
    assert ge["native_engine_handling"]["zero_optimization.stage3_max_live_parameters"].startswith("honoured")
    assert c.post("/api/v1/monitoring/create", json={"job_id": "jj", "window_size": 20}).json()["status"] == "created"
    # NaN in, JSON out (the reference answers 500 here)
    r = c.post("/api/v1/monitoring/ingest/single", content='{"job_id":"jj","step":1,"loss":NaN}',
               headers={"content-type": "application/json"})
    assert r.status_code == 200 and r.json()["healthy"] is False and r.json()["action"] == "rollback"
    assert c.get("/api/v1/monitoring/loss-curve/jj").status_code == 200
    assert c.delete("/api/v1/monitoring/reset/jj").json()["status"] == "reset"
    assert c.get("/api/v1/monitoring/jobs").status_code == 200        # best_loss=inf after reset
    assert c.get("/api/v1/monitoring/summary/none").status_code == 404
    t = c.get("/api/v1/nvlink/topology?mock=true").json()
    assert t["status"] == "success" and t["topology_matrix"]["GPU_0"]["GPU_1"] == "NV18"
    assert "b200zero_gpus_total" in c.get("/metrics").text
    assert c.get("/api/v1/jobs/unknown").status_code == 404
