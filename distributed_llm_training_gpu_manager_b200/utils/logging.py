"""Structured (JSON-lines) logging for workers and the API process. The reference has no logging at
all — only three ``print`` calls in a stub."""
from __future__ import annotations

import json
import logging
import os
import sys
import time
from typing import Any, Optional


class JsonFormatter(logging.Formatter):
    def format(self, record: logging.LogRecord) -> str:
        doc = {"ts": round(time.time(), 3), "level": record.levelname, "logger": record.name,
               "msg": record.getMessage(), "rank": int(os.environ.get("RANK", "0")),
               "job": os.environ.get("B200ZERO_JOB_ID")}
        extra = getattr(record, "fields", None)
        if isinstance(extra, dict):
            doc.update(extra)
        if record.exc_info:
            doc["exc"] = self.formatException(record.exc_info)
        return json.dumps(doc, default=str)


def get_logger(name: str = "b200zero", level: Optional[str] = None, stream=None) -> logging.Logger:
    log = logging.getLogger(name)
    if not log.handlers:
        h = logging.StreamHandler(stream or sys.stderr)
        h.setFormatter(JsonFormatter())
        log.addHandler(h)
        log.propagate = False
    log.setLevel((level or os.environ.get("B200ZERO_LOG", "INFO")).upper())
    return log


def log_event(log: logging.Logger, msg: str, **fields: Any) -> None:
    log.info(msg, extra={"fields": fields})
