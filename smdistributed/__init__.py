"""Import-path compatibility: ``import smdistributed.modelparallel.torch as smp`` resolves to
``smdistributed_modelparallel_b200`` (see smdistributed/modelparallel/torch/__init__.py)."""
