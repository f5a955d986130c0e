"""areal_vllm_b200 — B200-native (sm_100a) kernels for AReaL's per-token RL math hot path.

Drop-in modules (same names/signatures as the reference call sites import):

* ``areal_vllm_b200.functional``      <- ``areal.utils.functional``
* ``areal_vllm_b200.actor``           <- ``AdvNorm`` / ``grpo_loss_fn`` of ``areal.engine.ppo.actor``
* ``areal_vllm_b200.ppo_functional``  <- GAE entry points of ``realhf.impl.model.utils.ppo_functional``
                                         and the ``cugae`` binding
* ``areal_vllm_b200.flatten_param``   <- ``slice_intervals`` / ``set_intervals`` / ``merge_intervals``
* ``areal_vllm_b200.lm_engine``       <- ``compute_packed_sft_loss`` of ``areal.engine.sft.lm_engine``
* ``areal_vllm_b200.mcore_functional``<- the tensor-parallel branch of ``areal.experimental.megatron_actor``

Importing the package loads ``libb200rl.so``; it raises if the library was not built.
"""

from . import _lib  # noqa: F401  (fails loudly when the CUDA library is missing)
from . import ops  # noqa: F401
from . import functional, ppo_functional, flatten_param, actor, dist, lm_engine, mcore_functional  # noqa: F401

__version__ = "0.1.0"
