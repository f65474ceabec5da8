"""Engine library of the orchestrator: launcher, GPU fleet manager, loss monitor and spot
resiliency. Export list mirrors the reference (``ai_engine/__init__.py:9-17``) and adds the
types the reference's routers import by path."""
from .deepspeed_launcher import (DeepSpeedConfig, DeepSpeedLauncher, JobRecord, JobRegistry, LaunchResult,
                                 OffloadDevice, ZeROStage)
from .gpu_manager import (FleetSampler, GPUDevice, GPUFleetStatus, GPUHealthStatus, GPUManager, GPUProcess,
                          HealthThresholds)
from .loss_monitor import (AlertSeverity, HaltAction, HaltPolicy, LossSpikeMonitor, MonitorConfig, MonitorState,
                           SpikeAlert, TrainingMetrics)
from .spot_resiliency import PreemptionNotice, SpotInstanceResiliencyManager

__all__ = [
    "GPUManager", "GPUDevice", "GPUFleetStatus",
    "DeepSpeedLauncher", "DeepSpeedConfig",
    "LossSpikeMonitor", "TrainingMetrics", "SpikeAlert",
    # importable in the reference but not exported there
    "ZeROStage", "OffloadDevice", "LaunchResult", "MonitorConfig", "MonitorState", "AlertSeverity",
    "GPUProcess", "GPUHealthStatus", "SpotInstanceResiliencyManager",
    # new
    "JobRegistry", "JobRecord", "HealthThresholds", "FleetSampler", "HaltPolicy", "HaltAction",
    "PreemptionNotice",
]
