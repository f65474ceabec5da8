"""Hugging Face Llama layout <-> ours. transformers' own LlamaForCausalLM (random tiny config, fp32, CPU) is an
independent oracle for the model math: same logits from our nn.Module, same loss from the native executor."""
import pytest
import torch

transformers = pytest.importorskip("transformers")


def _tiny_hf(tie=False):
    from transformers import LlamaConfig, LlamaForCausalLM
    hc = LlamaConfig(vocab_size=1000, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                     num_attention_heads=2, num_key_value_heads=1, max_position_embeddings=256, rms_norm_eps=1e-5,
                     rope_theta=500000.0, tie_word_embeddings=tie, attention_bias=False, mlp_bias=False)
    torch.manual_seed(0)
    return hc, LlamaForCausalLM(hc).float().eval()


@pytest.mark.parametrize("tie", [False, True])
def test_module_matches_transformers_logits(tie):
    from distributed_llm_training_gpu_manager_b200.models import build_torch_model
    from distributed_llm_training_gpu_manager_b200.models.hf import config_from_hf, from_hf_state_dict, to_hf_state_dict
    hc, hf = _tiny_hf(tie)
    cfg = config_from_hf(hc)
    assert cfg.padded_vocab == 1024 and cfg.head_dim == 128 and cfg.n_kv_heads == 1
    named = from_hf_state_dict(hf.state_dict(), cfg)
    ours = build_torch_model(cfg).float().eval()
    ours.load_state_dict(named, strict=not tie)
    ids = torch.randint(0, 1000, (2, 64))
    with torch.no_grad():
        lh, lo = hf(input_ids=ids).logits, ours(ids)
    assert (lh - lo[..., :1000]).abs().max().item() < 1e-4
    back = to_hf_state_dict(named, cfg)
    for k, v in hf.state_dict().items():
        if k in back:
            assert torch.equal(back[k], v.float()), k


def test_native_engine_loads_hf_weights_and_matches_loss():
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.models.hf import config_from_hf, from_hf_state_dict
    from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine
    hc, hf = _tiny_hf()
    cfg = config_from_hf(hc)
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-4, warmup_steps=1,
        wall_clock_breakdown=False))
    eng = NativeLlamaEngine(cfg, ds, seq_len=64, device=torch.device("cpu"), seed=5)
    assert eng.load_full_params(from_hf_state_dict(hf.state_dict(), cfg)) == []
    ids = torch.randint(0, 1000, (2, 64))
    labels = torch.randint(0, 1000, (2, 64))
    with torch.no_grad():
        ref = torch.nn.functional.cross_entropy(hf(input_ids=ids).logits.view(-1, 1000), labels.view(-1)).item()
    loss = float(eng.forward(ids, labels))
    # the padded vocabulary rows are zero: they add exp(0)·24 to every partition function, a known offset
    with torch.no_grad():
        z = hf(input_ids=ids).logits.view(-1, 1000)
        lse_pad = torch.logsumexp(torch.cat([z, torch.zeros(z.shape[0], 24)], 1), 1)
        ref_padded = (lse_pad - z.gather(1, labels.view(-1, 1)).squeeze(1)).mean().item()
    assert loss == pytest.approx(ref_padded, rel=1e-4), (loss, ref, ref_padded)
    with pytest.raises(ValueError):
        eng.load_full_params({"embed.weight": torch.zeros(3, 3)}, strict=False)


def test_native_backward_matches_transformers_autograd():
    """Hand-written backward of the native executor (CPU restatements of the kernels) against autograd through
    transformers' implementation: every gradient, with a vocabulary that needs no padding so the losses are equal."""
    from transformers import LlamaConfig, LlamaForCausalLM
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.models.hf import config_from_hf, from_hf_state_dict
    from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine
    hc = LlamaConfig(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                     num_attention_heads=2, num_key_value_heads=1, max_position_embeddings=256, rms_norm_eps=1e-5,
                     rope_theta=500000.0, tie_word_embeddings=False, attention_bias=False, mlp_bias=False)
    torch.manual_seed(1)
    hf = LlamaForCausalLM(hc).float()
    cfg = config_from_hf(hc)
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-4, warmup_steps=1,
        gradient_clipping=0.0, wall_clock_breakdown=False))
    eng = NativeLlamaEngine(cfg, ds, seq_len=64, device=torch.device("cpu"), seed=5)
    eng.load_full_params(from_hf_state_dict(hf.state_dict(), cfg))
    ids = torch.randint(0, 1024, (2, 64))
    labels = torch.randint(0, 1024, (2, 64))
    ref_loss = torch.nn.functional.cross_entropy(hf(input_ids=ids).logits.view(-1, 1024), labels.view(-1))
    ref_loss.backward()
    loss = eng.forward(ids, labels)
    assert float(loss) == pytest.approx(ref_loss.item(), rel=1e-5)
    eng.backward(loss)
    sd = dict(hf.named_parameters())
    g = lambda k: sd[k].grad  # noqa: E731

    def close(ours, ref, name):
        scale = ref.abs().max().item() + 1e-12
        assert (ours.float() - ref).abs().max().item() / scale < 2e-4, name
    close(eng.grad_view("embed.weight"), g("model.embed_tokens.weight"), "embed")
    close(eng.grad_view("lm_head.weight"), g("lm_head.weight"), "lm_head")
    close(eng.grad_view("final_norm.weight"), g("model.norm.weight"), "final_norm")
    for i in range(2):
        h, p = f"model.layers.{i}.", f"blocks.{i}."
        close(eng.grad_view(p + "wqkv.weight"),
              torch.cat([g(h + "self_attn.q_proj.weight"), g(h + "self_attn.k_proj.weight"),
                         g(h + "self_attn.v_proj.weight")], 0), p + "wqkv")
        close(eng.grad_view(p + "wo.weight"), g(h + "self_attn.o_proj.weight"), p + "wo")
        close(eng.grad_view(p + "w_gate_up.weight"),
              torch.cat([g(h + "mlp.gate_proj.weight"), g(h + "mlp.up_proj.weight")], 0), p + "w_gate_up")
        close(eng.grad_view(p + "w_down.weight"), g(h + "mlp.down_proj.weight"), p + "w_down")
        close(eng.grad_view(p + "attn_norm.weight"), g(h + "input_layernorm.weight"), p + "attn_norm")
        close(eng.grad_view(p + "ffn_norm.weight"), g(h + "post_attention_layernorm.weight"), p + "ffn_norm")


def test_gpt2_module_matches_transformers():
    from transformers import GPT2Config, GPT2LMHeadModel
    from distributed_llm_training_gpu_manager_b200.models import build_torch_model
    from distributed_llm_training_gpu_manager_b200.models.hf import gpt2_config_from_hf, gpt2_from_hf_state_dict
    hc = GPT2Config(vocab_size=512, n_positions=128, n_embd=128, n_layer=2, n_head=4, activation_function="gelu_new",
                    resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
    torch.manual_seed(0)
    hf = GPT2LMHeadModel(hc).float().eval()
    cfg = gpt2_config_from_hf(hc)
    ours = build_torch_model(cfg).float().eval()
    ours.load_state_dict(gpt2_from_hf_state_dict(hf.state_dict(), cfg), strict=True)
    ids = torch.randint(0, 512, (2, 48))
    with torch.no_grad():
        lh, lo = hf(input_ids=ids).logits, ours(ids)
    assert (lh - lo[..., :512]).abs().max().item() < 1e-4
    # and the loss / gradients through both
    hf.train(); ours.train()
    labels = torch.randint(0, 512, (2, 48))
    l_ref = torch.nn.functional.cross_entropy(hf(input_ids=ids).logits.view(-1, 512), labels.view(-1))
    l_ref.backward()
    l_ours = ours(ids, labels)
    l_ours.backward()
    assert l_ours.item() == pytest.approx(l_ref.item(), rel=1e-5)
    ref_g = hf.transformer.h[0].mlp.c_fc.weight.grad.t()
    assert torch.allclose(ours.blocks[0].mlp_fc.weight.grad, ref_g, rtol=1e-3, atol=1e-6)


def test_trained_weights_export_to_transformers():
    """Train a few steps with the native engine, export with full_state_dict + to_hf_state_dict, load into a fresh
    transformers model: its logits must equal our module model's logits on the same weights."""
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.models import build_torch_model
    from distributed_llm_training_gpu_manager_b200.models.hf import config_from_hf, to_hf_state_dict
    from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine
    hc, hf = _tiny_hf()
    cfg = config_from_hf(hc)
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
        wall_clock_breakdown=False))
    eng = NativeLlamaEngine(cfg, ds, seq_len=32, device=torch.device("cpu"), seed=2)
    g = torch.Generator().manual_seed(0)
    for _ in range(3):
        ids = torch.randint(0, 1000, (2, 32), generator=g)
        eng.train_step(ids, ids)
    named = eng.full_state_dict()
    hf.load_state_dict(to_hf_state_dict(named, cfg), strict=True)
    ours = build_torch_model(cfg).float().eval()
    ours.load_state_dict(named, strict=True)
    ids = torch.randint(0, 1000, (2, 32), generator=g)
    with torch.no_grad():
        assert (hf(input_ids=ids).logits - ours(ids)[..., :1000]).abs().max().item() < 1e-4


def test_worker_starts_from_a_hugging_face_checkpoint(tmp_path):
    """`--init_hf DIR`: architecture and weights come from a transformers checkpoint directory; the first reported
    loss equals the loss transformers computes on the worker's first (synthetic) batch with those weights."""
    import json
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.engine import worker
    from distributed_llm_training_gpu_manager_b200.engine.trainer import synthetic_batches
    hc, hf = _tiny_hf()
    hf.save_pretrained(tmp_path / "ckpt", safe_serialization=True)
    launcher = DeepSpeedLauncher()
    cfg_path = launcher.write_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-4, warmup_steps=1,
        wall_clock_breakdown=False), str(tmp_path / "ds.json"))
    out = tmp_path / "result.json"
    assert worker.main(["--deepspeed_config", cfg_path, "--init_hf", str(tmp_path / "ckpt"), "--engine", "native",
                        "--steps", "2", "--seq_len", "32", "--result", str(out)]) == 0
    res = json.loads(out.read_text())
    ids, labels = synthetic_batches(1000, 2, 32, "cpu", 0, 1, seed=1234)(0)
    with torch.no_grad():
        z = hf(input_ids=ids).logits.view(-1, 1000)
        lse = torch.logsumexp(torch.cat([z, torch.zeros(z.shape[0], 24)], 1), 1)      # 24 zero rows of vocabulary padding
        ref = (lse - z.gather(1, labels.view(-1, 1)).squeeze(1)).mean().item()
    assert res["losses"][0] == pytest.approx(ref, rel=1e-4)


def test_checkpoint_consolidates_into_transformers_layout(tmp_path):
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.engine import checkpoint as ckpt, zero_to_fp32
    from distributed_llm_training_gpu_manager_b200.models.hf import config_from_hf
    from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine
    hc, hf = _tiny_hf()
    hc.save_pretrained(tmp_path / "cfg")
    cfg = config_from_hf(hc)
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
        wall_clock_breakdown=False))
    eng = NativeLlamaEngine(cfg, ds, seq_len=32, device=torch.device("cpu"), seed=2)
    ids = torch.randint(0, 1000, (2, 32))
    eng.train_step(ids, ids)
    eng.save_checkpoint(str(tmp_path / "ck"), tag="s1")
    ckpt.wait_for_pending_writes()
    assert zero_to_fp32.main([str(tmp_path / "ck"), str(tmp_path / "hf.pt"), "--hf_config", str(tmp_path / "cfg" / "config.json")]) == 0
    hf.load_state_dict(torch.load(tmp_path / "hf.pt"), strict=True)
    with torch.no_grad():
        z = hf(input_ids=ids).logits.view(-1, 1000)
    # same weights -> the engine's validation loss equals transformers' loss up to the 24 padded vocabulary rows
    with torch.no_grad():
        pad_logits = torch.nn.functional.linear(hf.model(input_ids=ids).last_hidden_state.view(-1, 256),
                                                eng.full_param("lm_head.weight")[1000:])
    lse_full = torch.logsumexp(torch.cat([z, pad_logits], 1), 1)
    ref = (lse_full - z.gather(1, ids.view(-1, 1)).squeeze(1)).mean().item()
    assert float(eng.eval_loss(ids, ids)) == pytest.approx(ref, rel=1e-4)
