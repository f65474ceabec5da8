"""NVLink / NVSwitch topology endpoint. The reference returns a hard-coded H100 matrix from
a router it never mounts (``backend/routers/nvlink.py:6-27``); this one reports the real
fabric through NVML (``parallel/topology.py``) and is mounted at ``/api/v1/nvlink``."""
from typing import Any, Dict

from fastapi import APIRouter

from ...parallel.topology import discover_topology

router = APIRouter()


@router.get("/topology")
async def get_nvlink_topology(mock: bool = False) -> Dict[str, Any]:
    topo = discover_topology(force_mock=mock)
    return {"status": "success", "node_type": topo.node_type, "interconnect": topo.interconnect,
            "bottlenecks": topo.bottlenecks(), "topology_matrix": topo.matrix_labels(),
            "num_gpus": topo.num_gpus, "links_per_gpu": topo.links_per_gpu, "source": topo.source,
            "recommended_rank_order": topo.rank_order()}
