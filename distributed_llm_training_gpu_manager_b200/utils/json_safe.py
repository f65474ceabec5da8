"""JSON that survives NaN/Inf. The reference's monitoring routes answer HTTP 500 whenever a
payload contains a non-finite float (NaN loss, ``best_loss=inf`` after reset);
here non-finite floats become ``null`` (or strings, if ``stringify``) before encoding."""
from __future__ import annotations

import math
from enum import Enum
from typing import Any

from fastapi.responses import JSONResponse


def sanitize(obj: Any, stringify: bool = False) -> Any:
    if isinstance(obj, float):
        if math.isfinite(obj):
            return obj
        return (("nan" if math.isnan(obj) else ("inf" if obj > 0 else "-inf")) if stringify else None)
    if isinstance(obj, dict):
        return {k: sanitize(v, stringify) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [sanitize(v, stringify) for v in obj]
    if isinstance(obj, Enum):
        return obj.value
    if hasattr(obj, "model_dump"):
        return sanitize(obj.model_dump(), stringify)
    return obj


class SafeJSONResponse(JSONResponse):
    def render(self, content: Any) -> bytes:
        return super().render(sanitize(content))
