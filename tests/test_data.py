"""Token-file input pipeline: deterministic, rank-disjoint, replayable batches with background prefetch."""
import numpy as np
import pytest
import torch

from distributed_llm_training_gpu_manager_b200.engine.data import TokenFileDataset, token_batches


def _write(tmp_path, n=10_000, dtype=np.uint16):
    arr = (np.arange(n) % 500).astype(dtype)
    p = tmp_path / "tokens.bin"
    arr.tofile(p)
    return p, arr


def test_dataset_windows(tmp_path):
    p, arr = _write(tmp_path)
    ds = TokenFileDataset(p, seq_len=64, vocab_size=512)
    assert ds.dtype == np.uint16 and len(ds) == (len(arr) - 1) // 64
    w = ds.window(3)
    assert w.shape == (65,) and np.array_equal(w, arr[192:257])
    with pytest.raises(ValueError):
        TokenFileDataset(p, seq_len=20_000)
    with pytest.raises(ValueError):
        TokenFileDataset(p, seq_len=64, dtype="float32")


def test_batches_are_shifted_deterministic_and_rank_disjoint(tmp_path):
    arr = np.arange(10_000, dtype=np.uint32)                                    # every window is unique
    p = tmp_path / "unique.bin"
    arr.tofile(p)
    mk = lambda r: token_batches(p, 4, 32, "cpu", rank=r, world_size=2, seed=7, dtype="uint32")  # noqa: E731
    a, b = mk(0), mk(1)
    starts = []
    for step in range(6):
        for g in (a, b):
            ids, labels = g(step)
            assert ids.shape == (4, 32) and ids.dtype == torch.int64
            assert torch.equal(ids[:, 1:], labels[:, :-1])                      # labels are the next tokens
            assert torch.equal(labels[:, -1], ids[:, -1] + 1)
            assert all(int(s) % 32 == 0 for s in ids[:, 0])                     # windows start on window boundaries
            starts += ids[:, 0].tolist()
    assert len(set(starts)) == 6 * 2 * 4                                        # no window served twice within the epoch
    # replay (rollback / resume): same micro-step -> same batch, from a fresh pipeline and from the same one
    ids0, _ = a(2)
    ids1, _ = mk(0)(2)
    assert torch.equal(ids0, ids1)
    a.close(); b.close()


def test_prefetch_matches_direct_assembly(tmp_path):
    p, _ = _write(tmp_path)
    g = token_batches(p, 2, 16, "cpu", seed=1, vocab_size=512, prefetch=3)
    ref = token_batches(p, 2, 16, "cpu", seed=1, vocab_size=512, prefetch=0)
    import time
    first = g(0)
    time.sleep(0.3)                                                             # let the prefetcher run ahead
    for step in range(1, 8):
        x, y = g(step)
        rx, ry = ref(step)
        assert torch.equal(x, rx) and torch.equal(y, ry)
    assert torch.equal(first[0], ref(0)[0])
    # epochs wrap with a new permutation
    spe = g.steps_per_epoch
    assert not torch.equal(g(0)[0], g(spe)[0])
    g.close(); ref.close()


def test_out_of_vocabulary_token_is_reported(tmp_path):
    p, _ = _write(tmp_path)
    g = token_batches(p, 2, 16, "cpu", vocab_size=100, prefetch=0)
    with pytest.raises(ValueError):
        g(0)


def test_trainer_on_token_file(tmp_path):
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.engine import initialize
    from distributed_llm_training_gpu_manager_b200.engine.trainer import Trainer, TrainerConfig
    from distributed_llm_training_gpu_manager_b200.models import build_torch_model, get_config
    mc = get_config("gpt2-tiny")
    # a learnable stream: next token = (token * 5 + 7) mod vocab
    toks = [3]
    for _ in range(20_000):
        toks.append((toks[-1] * 5 + 7) % mc.vocab_size)
    p = tmp_path / "stream.bin"
    np.asarray(toks, dtype=np.uint16).tofile(p)
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=1, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=4, gradient_accumulation_steps=1, learning_rate=3e-3, warmup_steps=2,
        wall_clock_breakdown=False))
    torch.manual_seed(0)
    eng = initialize(build_torch_model(mc), ds, device=torch.device("cpu"))
    batches = token_batches(p, 4, 32, "cpu", seed=3, vocab_size=mc.vocab_size)
    res = Trainer(eng, batches, TrainerConfig(total_steps=30, log_every=1000)).run()
    batches.close()
    assert res.status == "finished"
    losses = [h["loss"] for h in res.history]
    assert losses[-1] < losses[0] * 0.8


def test_trainer_validation_loss(tmp_path):
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.engine import initialize
    from distributed_llm_training_gpu_manager_b200.engine.trainer import Trainer, TrainerConfig, synthetic_batches
    from distributed_llm_training_gpu_manager_b200.models import build_torch_model, get_config
    mc = get_config("gpt2-tiny")
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=1, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=4, gradient_accumulation_steps=1, learning_rate=3e-3, warmup_steps=2,
        wall_clock_breakdown=False))
    torch.manual_seed(0)
    eng = initialize(build_torch_model(mc), ds, device=torch.device("cpu"))
    train = synthetic_batches(mc.vocab_size, 4, 32, "cpu", seed=1)
    val = synthetic_batches(mc.vocab_size, 4, 32, "cpu", seed=99)
    res = Trainer(eng, train, TrainerConfig(total_steps=12, log_every=1000, eval_batches=val, eval_every=4,
                                            eval_steps=2)).run()
    vals = [(h["step"], h["val_loss"]) for h in res.history if "val_loss" in h]
    assert [s for s, _ in vals] == [4, 8, 12]
    assert vals[-1][1] < vals[0][1]                                              # the learnable stream generalises
