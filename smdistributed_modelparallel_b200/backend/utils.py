"""Small backend helpers (reference smdistributed/modelparallel/backend/utils.py:80-149)."""

from __future__ import annotations

import json
import math
import os


def bijection_2d(a: int, b: int) -> int:
    """Cantor pairing N x N -> N (unique ids from (peer, counter) pairs)."""
    return (a + b) * (a + b + 1) // 2 + b


def inverse_bijection_2d(z: int) -> tuple[int, int]:
    w = (math.isqrt(8 * z + 1) - 1) // 2
    b = z - w * (w + 1) // 2
    return w - b, b


def get_divisibility_error_str(what: str, value: int, by_what: str, by: int) -> str:
    return f"{what} ({value}) must be divisible by {by_what} ({by})."


def upload_metrics_to_studio(metrics: dict) -> None:
    """Persist one-time job metrics.  The reference posts them to SageMaker Studio; outside SageMaker
    (always, here) they are written as JSON when ``SMP_METRICS_FILE`` is set, otherwise dropped."""
    path = os.environ.get("SMP_METRICS_FILE")
    if not path:
        return
    try:
        with open(path, "w") as f:
            json.dump(metrics, f, indent=1, default=str)
    except OSError:
        pass
