"""Namespace shim; the library lives in ``smdistributed_modelparallel_b200``."""
