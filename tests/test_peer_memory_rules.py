"""Source-level guard on how kernels may touch another GPU's memory.

Rule: TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load
from another GPU's memory hung the GPU. The kernels go further: no tensor map, no bulk copy and no
library GEMM operand is ever built over a peer's buffer — remote rows arrive through plain
``ld.global`` and are written to local memory first. These tests fail if a change breaks that."""
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
PKG = ROOT / "distributed_llm_training_gpu_manager_b200"


def _sources():
    return sorted(list(CSRC.rglob("*.cu")) + list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.cpp")) +
                  list(CSRC.rglob("*.h")))


def _strip_comments(text: str) -> str:
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"//[^\n]*", "", text)


def test_no_multicast_bulk_loads_in_csrc():
    """No `cp.async.bulk…` instruction with `.multicast::cluster` anywhere in csrc/ (the only
    multicast instruction is `tcgen05.commit`, which signals barriers and moves no data)."""
    bad = []
    for f in _sources():
        code = _strip_comments(f.read_text())
        for m in re.finditer(r"cp\.async\.bulk[^\"\n;]*multicast", code):
            bad.append((f.name, m.group(0)))
    assert not bad, bad
    for f in _sources():
        code = _strip_comments(f.read_text())
        for m in re.finditer(r"[\w.:]*\.multicast::cluster[\w.:]*", code):
            assert m.group(0).startswith("tcgen05.commit"), (f.name, m.group(0))


# Names that hold another rank's memory in the host launchers (pointer tables filled from IPC / symmetric handles).
_PEER_NAMES = re.compile(r"\b(shard_ptrs|grad_ptrs|param_ptrs|signal_ptrs|data_ptrs|peer_ptrs|grad_mc|mc_ptr|pt\.data|"
                         r"p\.shard|P\.shard|shard\[)")


def test_tensor_maps_cover_local_buffers_only():
    """Every `make_tmap_2d_bf16(&map, base, …)` call takes a base that is a plain kernel argument of the local
    device (activations, the local gathered weight, gradients) — never an entry of a peer pointer table."""
    calls = []
    for f in _sources():
        code = _strip_comments(f.read_text())
        for m in re.finditer(r"make_tmap_2d_bf16\(\s*&\w+\s*,\s*([^,]+),", code):
            calls.append((f.name, m.group(1).strip()))
    assert len(calls) >= 10, "expected the GEMM / attention / AG-GEMM launchers to be scanned"
    for name, base in calls:
        assert not _PEER_NAMES.search(base), f"{name}: tensor map over peer memory: {base}"
        assert re.fullmatch(r"[A-Za-z_]\w*", base) or base == "const void* base", (name, base)


def test_python_never_hands_peer_pointers_to_tma_or_library_gemms():
    """`_ptrs_for()` / `ptr_array()` (tables of every rank's buffer) may only flow into the fused
    collective entry points and the AG→GEMM wrappers, whose kernels read peers with `ld.global`."""
    for f in sorted(PKG.rglob("*.py")):
        text = f.read_text()
        for m in re.finditer(r"^.*(_ptrs_for\(|\.ptr_array\(\)).*$", text, flags=re.M):
            line = m.group(0)
            if line.lstrip().startswith(("def ", "#")):
                continue
            # the table must not be passed to the plain GEMM / attention wrappers or to torch
            assert not re.search(r"bz_gemm_bf16|bz_flash_attn|torch\.(mm|matmul|addmm)|F\.linear", line), (f.name, line)
