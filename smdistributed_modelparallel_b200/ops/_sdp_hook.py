"""Registry connecting the sharded-data-parallel engine to the GEMM front-end: gathers deferred to the GEMM
that first reads a unit's weight, keyed by the weight's device address (see parallel/sdp_fused.py)."""

pending_gathers: dict = {}
