"""Regression tests for the round-1 review findings that need no GPU: collective preemption, back-to-back
asynchronous saves, vector-granular segment tables."""
import os
import threading
from pathlib import Path
import time

import pytest
import torch

from distributed_llm_training_gpu_manager_b200.engine import checkpoint as ckpt
from distributed_llm_training_gpu_manager_b200.engine import initialize
from distributed_llm_training_gpu_manager_b200.engine.partition import build_layout
from distributed_llm_training_gpu_manager_b200.engine.trainer import Trainer, TrainerConfig, synthetic_batches
from distributed_llm_training_gpu_manager_b200.parallel.fused import seg_len

from dist_utils import run_ranks
from test_zero_engine import MB, SEQ, VOCAB, _model, ds_dict


# ---------------------------------------------------------------- preemption is agreed between ranks
def _preempt_rank(rank, world, tmp):
    eng = initialize(_model(), ds_dict(2), device="cpu")
    data = synthetic_batches(VOCAB, MB, SEQ, "cpu", rank, world)
    tr = Trainer(eng, data, TrainerConfig(total_steps=12, save_dir=tmp, save_every=0, async_save=False, log_every=100))
    orig = eng._optimizer_step

    def step_with_signal():
        # only rank 1 is told to stop, and only from its 4th step on (a SIGTERM seen on one side of a boundary)
        if rank == 1 and eng.global_steps == 3:
            tr._on_sigterm(15, None)
        return orig()
    eng._optimizer_step = step_with_signal
    res = tr.run()
    return res.status, res.final_step


def test_preemption_bit_is_agreed_by_all_ranks(tmp_path):
    out = run_ranks(_preempt_rank, 2, (str(tmp_path),))
    assert out[0] == out[1] == ("preempted", 4)          # both ranks stop after the same step
    assert (tmp_path / "preempt_step4" / "manifest.json").exists()
    assert ckpt.verify_checkpoint(str(tmp_path), "preempt_step4") == (True, "ok")


def test_single_rank_preempt_request_reaches_metrics():
    eng = initialize(_model(), ds_dict(0), device="cpu")
    data = synthetic_batches(VOCAB, MB, SEQ, "cpu")
    eng.backward(eng(*data(0)))
    assert eng.step()["preempt"] is False
    eng.preempt_requested = True
    eng.backward(eng(*data(1)))
    assert eng.step()["preempt"] is True


# ---------------------------------------------------------------- a second save must not tear the first
def test_async_save_followed_by_second_save_keeps_first_intact(tmp_path, monkeypatch):
    eng = initialize(_model(), ds_dict(1), device="cpu")
    data = synthetic_batches(VOCAB, MB, SEQ, "cpu")
    eng.backward(eng(*data(0))); eng.step()
    want_master = eng.master.clone()
    want_scale = dict(eng.scaler.state_dict())
    slow = threading.Event()
    real_save = ckpt._atomic_torch_save

    def slow_save(obj, path):
        if not slow.is_set():
            slow.set()
            time.sleep(0.5)                # first writer is still serialising when the second save is requested
        real_save(obj, path)
    monkeypatch.setattr(ckpt, "_atomic_torch_save", slow_save)
    eng.save_checkpoint(str(tmp_path), tag="global_step1", async_write=True)
    eng.backward(eng(*data(1))); eng.step()
    eng.save_checkpoint(str(tmp_path), tag="global_step2", async_write=True)
    ckpt.wait_for_pending_writes()
    for tag in ("global_step1", "global_step2"):
        assert ckpt.verify_checkpoint(str(tmp_path), tag) == (True, "ok")
    sd = torch.load(tmp_path / "global_step1" / ckpt.optim_file(0, False), weights_only=False)["optimizer_state_dict"]
    assert torch.equal(sd["single_partition_of_fp32_groups"][0], want_master)
    assert sd["loss_scaler"] == want_scale
    assert sd["base_optimizer_state"]["state"][0]["step"] == 1


# ---------------------------------------------------------------- fused kernels move whole 16-byte vectors
def test_segment_lengths_are_whole_vectors():
    specs = [("w", (64, 48), 0), ("bias_odd", (13,), 0), ("vocab_odd", (50257,), 0), ("blk.w", (128, 64), 1)]
    L = build_layout(specs, ["root", "b0"], world_size=4, partition=True, persistence_threshold=1000)
    seen = {}
    for u in L.units:
        for s in u.params:
            n = seg_len(s, 4)
            seen[s.name] = n
            assert n % 8 == 0
            if s.partitioned:
                assert n == s.numel and (n // 4) % 8 == 0
            else:
                assert n >= s.numel
    assert seen["bias_odd"] == 16 and seen["vocab_odd"] == 50272
    # padded lengths never run into the next tensor of the unit or of the shard
    for u in L.units:
        ordered = sorted(u.params, key=lambda s: s.full_offset)
        for a, b in zip(ordered, ordered[1:]):
            assert a.full_offset + seg_len(a, 4) <= b.full_offset
        assert ordered[-1].full_offset + seg_len(ordered[-1], 4) <= u.full_numel
    repl = sorted((s for u in L.units for s in u.params if not s.partitioned), key=lambda s: s.shard_offset)
    for a, b in zip(repl, repl[1:]):
        assert a.shard_offset + seg_len(a, 4) <= b.shard_offset
    assert repl[-1].shard_offset + seg_len(repl[-1], 4) <= L.shard_numel


# ---------------------------------------------------------------- stage3_max_live_parameters drives the block buffers
def _slots_rank(rank, world, max_live, reuse):
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.models import get_config
    from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine
    mc = get_config("llama-tiny", n_layers=6)
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
        stage3_max_live_parameters=max_live, stage3_max_reuse_distance=reuse, stage3_param_persistence_threshold=1000,
        wall_clock_breakdown=False))
    eng = NativeLlamaEngine(mc, ds, seq_len=32, device=torch.device("cpu"), seed=3)
    gathers = []
    real = eng._gather_unit

    def counting(ui, stream=None):
        gathers.append(ui)
        return real(ui, stream=stream)
    eng._gather_unit = counting
    g = torch.Generator().manual_seed(100 + rank)
    losses = []
    for _ in range(3):
        ids = torch.randint(0, mc.vocab_size, (2, 32), generator=g)
        losses.append(eng.train_step(ids, ids)["loss"])
    blk = max(u.full_numel for u in eng.layout.units[1:])
    return losses, eng._n_pslots, len([u for u in gathers if u >= 1]), blk


def test_max_live_parameters_sets_resident_blocks_and_saves_gathers():
    blk = run_ranks(_slots_rank, 2, (1, 10 ** 9))[0][3]
    base = run_ranks(_slots_rank, 2, (2 * blk, 10 ** 9))           # room for two blocks: the pipeline's minimum
    wide = run_ranks(_slots_rank, 2, (4 * blk + 1, 10 ** 9))       # four blocks stay gathered
    near = run_ranks(_slots_rank, 2, (4 * blk + 1, 1))             # reuse distance too short to keep anything extra
    assert base[0][1] == 2 and wide[0][1] == 4 and near[0][1] == 2
    assert wide[0][0] == base[0][0] == near[0][0]                   # same arithmetic, different residency
    # 6 blocks, 3 steps: every block is gathered for forward; backward re-gathers all but the ones still alive
    assert base[0][2] == 3 * (6 + 4) and wide[0][2] == 3 * (6 + 2)


# ---------------------------------------------------------------- offload_param: host-resident 16-bit shards
def _poffload_rank(rank, world, offload_param, offload_opt, native):
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.models import get_config
    from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer=offload_opt, offload_param=offload_param, bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
        stage3_param_persistence_threshold=1000, wall_clock_breakdown=False))
    if native:
        mc = get_config("llama-tiny", n_layers=4)
        eng = NativeLlamaEngine(mc, ds, seq_len=32, device=torch.device("cpu"), seed=3)
        step = eng.train_step
        vocab = mc.vocab_size
    else:
        eng = initialize(_model(), ds, device="cpu")
        vocab = VOCAB

        def step(ids, lab):
            eng.backward(eng(ids, lab))
            return eng.step()
    g = torch.Generator().manual_seed(100 + rank)
    losses = []
    for _ in range(3):
        ids = torch.randint(0, vocab, (2, 32), generator=g)
        losses.append(step(ids, ids)["loss"])
    return losses, eng.param_offload, eng.shard_params.device.type, getattr(eng, "h2d_param_bytes", 0)


@pytest.mark.parametrize("native", [True, False])
@pytest.mark.parametrize("world", [1, 2])
def test_offload_param_matches_resident_shards(native, world):
    base = run_ranks(_poffload_rank, world, ("none", "none", native))
    off = run_ranks(_poffload_rank, world, ("cpu", "cpu", native))
    for (l0, on0, _, b0), (l1, on1, dev1, b1) in zip(base, off):
        assert not on0 and b0 == 0
        assert on1 and dev1 == "cpu" and b1 > 0          # every gather staged its rows from the host shard
        assert l1 == l0


# ---------------------------------------------------------------- descriptors travel between ranks (VMM handle exchange)
def _fd_rank(rank, world):
    import os
    import tempfile
    from distributed_llm_training_gpu_manager_b200.parallel.symm import exchange_fds
    f = tempfile.TemporaryFile()
    f.write(f"rank{rank}".encode()); f.flush()
    got = exchange_fds([f.fileno()], rank, world, tag="t")
    seen = []
    for r in range(world):
        fd = got[r][0]
        seen.append(os.pread(fd, 16, 0).decode())
        if r != rank:
            os.close(fd)
    return seen


@pytest.mark.parametrize("world", [2, 4])
def test_file_descriptors_are_exchanged_all_to_all(world):
    for seen in run_ranks(_fd_rank, world):
        assert seen == [f"rank{r}" for r in range(world)]


def test_segment_table_packs_partitioned_and_push_into_one_word():
    from distributed_llm_training_gpu_manager_b200.parallel.fused import segment_table
    t = segment_table([(0, 64, 0, True), (64, 16, 128, False, True), (80, 32, 256, True, True)], "cpu")
    assert t.shape == (3, 4) and t.dtype == torch.int64
    flags = t[:, 3].tolist()
    assert [f & 0xFFFFFFFF for f in flags] == [1, 0, 1]          # int32 partitioned (low half, little endian)
    assert [f >> 32 for f in flags] == [0, 1, 1]                  # int32 push (high half)


# ---------------------------------------------------------------- a job that stops without anybody dying
def test_launcher_detects_a_stalled_job_and_relaunches_it(tmp_path):
    """Rank 1 stops at the top of step 12 and never returns (first start only); rank 0 then waits for it inside its
    next collective. Nobody dies, so only the heartbeat can tell: the launcher stops the job after the stall window,
    relaunches it, the workers resume from the newest verified checkpoint and finish."""
    import json
    import subprocess
    import sys
    from dist_utils import free_port
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    cfg_path = tmp_path / "ds.json"
    cfg_path.write_text(json.dumps(DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=2, offload_optimizer="none", offload_param="none", fp16_enabled=False, bf16_enabled=False,
        train_micro_batch_size_per_gpu=4, gradient_accumulation_steps=1, learning_rate=3e-3, warmup_steps=5,
        total_training_steps=200, wall_clock_breakdown=False))))
    result = tmp_path / "result.json"
    env = dict(os.environ, B200ZERO_BACKEND="gloo", B200ZERO_HANG_AT="12:1", MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="2")
    res = subprocess.run(
        [sys.executable, "-m", "distributed_llm_training_gpu_manager_b200.engine.launch", "--num_gpus=2",
         f"--master_port={free_port()}", "--max_restarts=1", "--stall_timeout=8", "--stall_startup=240",
         "--placement=linear", "builtin:gpt2-tiny", f"--deepspeed_config={cfg_path}", "--steps=25",
         f"--save_dir={tmp_path / 'ckpt'}", "--save_every=5", f"--result={result}", "--seq_len=32", "--window=20"],
        env=env, capture_output=True, text=True, timeout=600, cwd=str(Path(__file__).resolve().parent.parent))
    log = res.stdout + res.stderr
    assert res.returncode == 0, log[-3000:]
    assert "no rank has completed a step" in log and "job stalled; restart 1/1" in log, log[-3000:]
    assert "on NCCL collectives" in log                      # the relaunch is asked to leave the fused collectives
    assert "resumed from global_step10" in log, log[-3000:]
    out = json.loads(result.read_text())
    assert out["status"] == "finished" and out["final_step"] == 25


# ---------------------------------------------------------------- emitted keys that used to be parsed and dropped
def _keys_rank(rank, world, overrides):
    d = ds_dict(2)
    d["zero_optimization"].update(overrides.get("zero", {}))
    d.update(overrides.get("top", {}))
    eng = initialize(_model(), d, device="cpu")
    data = synthetic_batches(VOCAB, MB, SEQ, "cpu", rank, world)
    losses = []
    for _ in range(4):
        eng.backward(eng(*data(eng.micro_steps)))
        losses.append(eng.step()["loss"])
    return losses, eng.config_notes


def test_reduce_scatter_false_and_prescale_are_honoured_and_equivalent():
    base = run_ranks(_keys_rank, 2, ({},))
    allred = run_ranks(_keys_rank, 2, ({"zero": {"reduce_scatter": False}},))
    presc = run_ranks(_keys_rank, 2, ({"top": {"prescale_gradients": True, "gradient_predivide_factor": 2.0}},))
    for (l0, n0), (l1, n1), (l2, _n2) in zip(base, allred, presc):
        assert l1 == pytest.approx(l0, rel=1e-5) and l2 == pytest.approx(l0, rel=1e-5)
        assert not n0 and any("reduce_scatter=false" in n for n in n1)


def test_config_audit_names_what_is_fixed_here():
    from distributed_llm_training_gpu_manager_b200.engine.config import audit, load_config
    d = ds_dict(2)
    d["zero_optimization"].update(reduce_bucket_size=1000, allgather_bucket_size=1000, contiguous_gradients=False)
    notes = audit(load_config(d), unit_numels=[5000, 70000], world_size=2)
    assert len(notes) == 3 and "reduce_bucket_size=1000" in notes[0] and "70000" in notes[0]
    assert "allgather_bucket_size" in notes[1] and "contiguous_gradients" in notes[2]
    assert audit(load_config(ds_dict(3)), unit_numels=[5000], world_size=2) == []


def test_capacity_planner_knows_parameter_offload_and_resident_blocks():
    from distributed_llm_training_gpu_manager_b200.models import get_config
    from distributed_llm_training_gpu_manager_b200.models.native_llama import static_bytes
    c70 = get_config("llama3-70b")
    base = static_bytes(c70, 8, offload_optimizer=True)
    both = static_bytes(c70, 8, offload_optimizer=True, offload_param=True)
    assert both["param_shard"] < 0.1 * base["param_shard"]            # only the staging buffers stay on the device
    assert both["host_pinned"] - base["host_pinned"] == pytest.approx(2.0 * c70.num_params() / 8, rel=0.02)
    wide = static_bytes(c70, 8, offload_optimizer=True, param_slots=4)
    assert wide["param_buffers"] > base["param_buffers"] and wide["device_total"] > base["device_total"]


def test_rank_core_sets_even_and_numa():
    from distributed_llm_training_gpu_manager_b200.engine.launch import _parse_cpulist, core_sets
    assert _parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert core_sets(4, "none") == [None] * 4
    even = core_sets(4, "even", usable=list(range(10)))
    assert even == [[0, 1], [2, 3], [4, 5], [6, 7]]                       # disjoint, contiguous, leftovers unused
    cpus = {0: list(range(0, 8)), 1: list(range(8, 16))}
    numa = core_sets(4, "numa", numa_of_rank=[0, 0, 1, 1], usable=list(range(16)), node_cpus=lambda n: cpus[n])
    assert numa == [[0, 1, 2, 3], [4, 5, 6, 7], [8, 9, 10, 11], [12, 13, 14, 15]]
    # unknown NUMA node of some GPU -> even split; more ranks than cores -> leave affinity alone
    assert core_sets(2, "numa", numa_of_rank=[0, None], usable=list(range(4))) == [[0, 1], [2, 3]]
    assert core_sets(8, "even", usable=[0, 1, 2]) == [None] * 8


def test_launcher_binds_ranks_to_disjoint_cores(tmp_path):
    import subprocess
    import sys
    if len(os.sched_getaffinity(0)) < 2:
        pytest.skip("needs two usable cores")
    script = tmp_path / "show.py"
    script.write_text("import os\nopen(os.environ['OUT'] + os.environ['RANK'], 'w').write(','.join(map(str, sorted(os.sched_getaffinity(0)))))\n")
    env = dict(os.environ, OUT=str(tmp_path / "aff"))
    res = subprocess.run([sys.executable, "-m", "distributed_llm_training_gpu_manager_b200.engine.launch", "--num_gpus=2",
                          "--placement=linear", "--bind_cores=even", str(script), "--deepspeed_config=x"],
                         env=env, capture_output=True, text=True, timeout=120,
                         cwd=str(Path(__file__).resolve().parent.parent))
    assert res.returncode == 0, res.stderr[-1500:]
    a = set((tmp_path / "aff0").read_text().split(","))
    b = set((tmp_path / "aff1").read_text().split(","))
    assert a and b and not (a & b)


def test_trainer_collects_garbage_on_a_common_schedule(monkeypatch):
    import gc
    eng = initialize(_model(), ds_dict(0), device="cpu")
    data = synthetic_batches(VOCAB, MB, SEQ, "cpu")
    calls, states = [], []
    real = gc.collect
    monkeypatch.setattr(gc, "collect", lambda *a: (calls.append(eng.global_steps), states.append(gc.isenabled()), real(*a))[2])
    res = Trainer(eng, data, TrainerConfig(total_steps=6, gc_every=2, log_every=100)).run()
    assert res.status == "finished" and res.final_step == 6
    assert calls == [0, 2, 4, 6] and not any(states[1:])        # automatic GC is off while the loop runs ...
    assert gc.isenabled()                                        # ... and back on afterwards


def test_every_key_the_launcher_can_emit_is_accounted_for():
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.engine.config import KEY_HANDLING, explain, key_paths
    L = DeepSpeedLauncher()
    cfgs = list(L.presets().values()) + list(L.b200_presets().values()) + [
        DeepSpeedConfig(), DeepSpeedConfig(zero_stage=0), DeepSpeedConfig(zero_stage=1, bf16_enabled=True, fp16_enabled=False),
        DeepSpeedConfig(zero_stage=2, offload_optimizer="nvme", elastic_training=True, activation_checkpointing=False)]
    seen = set()
    for c in cfgs:
        seen |= set(key_paths(L.generate_config(c)))
    assert len(seen) >= 55
    missing = sorted(seen - set(KEY_HANDLING))
    assert not missing, f"emitted but not described in engine/config.py KEY_HANDLING: {missing}"
    ex = explain(L.generate_config(DeepSpeedConfig()))
    assert ex["zero_optimization.offload_param.device"].startswith("honoured")
    assert all(v.split(":")[0].split()[0] in ("honoured", "fixed", "n/a") for v in ex.values())
