"""Construction-time interposition on ``nn.Module``.

The reference replaces ``nn.Module.__init__`` so that every module records the partition and
tensor-parallel context it was created in (reference smdistributed/modelparallel/torch/patches/
__init__.py:7-18, patch_manager.py:1-154) and additionally swaps every module class' ``forward`` for a
dispatcher.  Here only the constructor is interposed; remote execution under pipeline parallelism is
driven by forward hooks the executor installs on partition boundaries (``runtime/server.py``), so
unpartitioned models run their original ``forward`` untouched.
"""

from __future__ import annotations

import torch.nn as nn


class PatchManager:
    def __init__(self):
        self._orig_init = None

    def patch_constructor(self):
        if self._orig_init is not None:
            return
        orig = nn.Module.__init__
        self._orig_init = orig

        def recording_init(module, *args, **kwargs):
            orig(module, *args, **kwargs)
            from ..state_mod import state

            if getattr(state, "initialized", False) and state.module_manager is not None:
                state.module_manager.record_module_construction(module)

        nn.Module.__init__ = recording_init

    def restore(self):
        if self._orig_init is not None:
            nn.Module.__init__ = self._orig_init
            self._orig_init = None

    def reset(self):
        pass


patch_manager = PatchManager()
