"""NVLink / NVSwitch topology discovery and rank placement.

Replaces the reference's canned ``"HGX H100 8-GPU"`` dictionary
(``backend/routers/nvlink.py:13-27``) with what NVML actually reports: per-GPU link
state, the remote end of every link (a peer GPU's PCI bus id, or an NVSwitch), P2P
status between pairs and CPU/NUMA affinity. On an HGX B200 every GPU has 18 NVLink-5
links into the NVSwitch fabric, so all pairs are equidistant ("NV18") and any rank order
is as good as any other; on a partially-connected box :meth:`Topology.rank_order` orders
ranks so neighbours in the ring share the most links.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional


@dataclass
class Topology:
    num_gpus: int
    names: List[str]
    links_per_gpu: List[int]                     # active NVLink links per GPU
    direct: Dict[int, Dict[int, int]]            # direct GPU<->GPU link counts
    via_switch: List[int]                        # links per GPU that terminate on an NVSwitch
    numa: List[Optional[int]] = field(default_factory=list)
    pci: List[str] = field(default_factory=list)
    source: str = "nvml"

    @property
    def has_nvswitch(self) -> bool:
        return any(n > 0 for n in self.via_switch)

    @property
    def interconnect(self) -> str:
        if self.has_nvswitch:
            return "NVSwitch Fabric"
        if any(self.links_per_gpu):
            return "NVLink (direct)"
        return "PCIe"

    @property
    def node_type(self) -> str:
        if not self.num_gpus:
            return "no GPUs"
        base = self.names[0].replace("NVIDIA ", "")
        return f"HGX {base.split('-')[0].split()[0]} {self.num_gpus}-GPU" if self.has_nvswitch else f"{self.num_gpus}x {base}"

    def link_counts(self) -> Dict[int, Dict[int, int]]:
        """Effective links between every pair: through an NVSwitch a pair can use
        min(links_i, links_j); otherwise only direct links count."""
        out: Dict[int, Dict[int, int]] = {}
        for i in range(self.num_gpus):
            out[i] = {}
            for j in range(self.num_gpus):
                if i == j:
                    continue
                n = self.direct.get(i, {}).get(j, 0)
                if self.via_switch[i] and self.via_switch[j]:
                    n = max(n, min(self.via_switch[i], self.via_switch[j]))
                out[i][j] = n
        return out

    def matrix_labels(self) -> Dict[str, Dict[str, str]]:
        lc = self.link_counts()
        lab: Dict[str, Dict[str, str]] = {}
        for i in range(self.num_gpus):
            row = {}
            for j in range(self.num_gpus):
                if i == j:
                    continue
                n = lc[i][j]
                if n > 0:
                    row[f"GPU_{j}"] = f"NV{n}"
                else:
                    same = (self.numa and self.numa[i] is not None and self.numa[i] == self.numa[j])
                    row[f"GPU_{j}"] = "NODE" if same else "SYS"
            lab[f"GPU_{i}"] = row
        return lab

    def bottlenecks(self) -> List[Dict[str, Any]]:
        out: List[Dict[str, Any]] = []
        best = max(self.links_per_gpu, default=0)
        for i, n in enumerate(self.links_per_gpu):
            if best and n < best:
                out.append({"type": "NVLink", "location": f"GPU_{i}", "severity": "high" if n == 0 else "medium",
                            "detail": f"{n}/{best} links active"})
        lc = self.link_counts()
        for i in range(self.num_gpus):
            for j in range(i + 1, self.num_gpus):
                if lc[i][j] == 0:
                    out.append({"type": "PCIe", "location": f"GPU_{i} <-> GPU_{j}", "severity": "medium",
                                "detail": "no NVLink path; traffic crosses PCIe/host"})
        return out

    def rank_order(self) -> List[int]:
        """Greedy chain through the link graph (uniform fabric → identity)."""
        if self.num_gpus <= 2 or self.has_nvswitch:
            return list(range(self.num_gpus))
        lc = self.link_counts()
        order, left = [0], set(range(1, self.num_gpus))
        while left:
            cur = order[-1]
            nxt = max(left, key=lambda j: (lc[cur][j], -j))
            order.append(nxt)
            left.remove(nxt)
        return order


def mock_topology(num_gpus: int = 8, links: int = 18, name: str = "NVIDIA B200") -> Topology:
    return Topology(num_gpus=num_gpus, names=[name] * num_gpus, links_per_gpu=[links] * num_gpus, direct={},
                    via_switch=[links] * num_gpus, numa=[i * 2 // max(num_gpus, 1) for i in range(num_gpus)],
                    pci=[f"00000000:{0x10 + i:02X}:00.0" for i in range(num_gpus)], source="mock")


def discover_topology(force_mock: bool = False) -> Topology:
    """Query NVML; fall back to an empty topology (or a mock 8×B200 when asked)."""
    if force_mock or os.environ.get("B200ZERO_MOCK_TOPOLOGY") == "1":
        return mock_topology()
    try:
        import pynvml as nv
        nv.nvmlInit()
        n = nv.nvmlDeviceGetCount()
    except Exception:
        return Topology(0, [], [], {}, [], source="none")

    def s(v):
        return v.decode() if isinstance(v, bytes) else str(v)

    handles = [nv.nvmlDeviceGetHandleByIndex(i) for i in range(n)]
    pci = []
    for h in handles:
        try:
            pci.append(s(nv.nvmlDeviceGetPciInfo(h).busId).lower())
        except Exception:
            pci.append("")
    by_pci = {p: i for i, p in enumerate(pci) if p}
    names, links, via_switch, numa = [], [], [], []
    direct: Dict[int, Dict[int, int]] = {i: {} for i in range(n)}
    max_links = getattr(nv, "NVML_NVLINK_MAX_LINKS", 18)
    for i, h in enumerate(handles):
        try:
            names.append(s(nv.nvmlDeviceGetName(h)))
        except Exception:
            names.append(f"GPU {i}")
        up = sw = 0
        for link in range(max_links):
            try:
                if not nv.nvmlDeviceGetNvLinkState(h, link):
                    continue
            except Exception:
                break
            up += 1
            try:
                remote = nv.nvmlDeviceGetNvLinkRemotePciInfo(h, link)
                rid = s(remote.busId).lower()
            except Exception:
                rid = ""
            peer = by_pci.get(rid)
            if peer is not None and peer != i:
                direct[i][peer] = direct[i].get(peer, 0) + 1
            else:
                sw += 1  # remote end is not one of our GPUs: an NVSwitch port
        links.append(up)
        via_switch.append(sw)
        try:
            numa.append(nv.nvmlDeviceGetNumaNodeId(h))
        except Exception:
            numa.append(None)
    return Topology(n, names, links, direct, via_switch, numa, pci, source="nvml")


def rank_to_device(local_rank: int, topo: Optional[Topology] = None) -> int:
    """CUDA device index for a local rank, honouring the topology-derived order."""
    topo = topo or discover_topology()
    order = topo.rank_order()
    return order[local_rank % len(order)] if order else local_rank
