"""evo_b200 -- B200-native E-step / M-step of tvlearn/evo's BSC and ES3C models.

Same model / K-state-container / driver-loop API as the reference (`evo.models`, `evo.variational`);
the compute runs in hand-written sm_100a CUDA kernels behind the C-ABI of libevo_b200.so.
There is no CPU fallback: importing the models without the built library raises.
"""
from . import _lib  # noqa: F401
from .comm import Comm, shard_bounds, split_rows  # noqa: F401


def __getattr__(name):
    # models need torch + the CUDA library; keep `import evo_b200` light for the CPU-only tests
    if name == "BSC":
        from .bsc import BSC
        return BSC
    if name == "SSSC":
        from .sssc import SSSC
        return SSSC
    if name in ("init_states", "vary_Kn", "KStates"):
        from . import variational
        return getattr(variational, name)
    raise AttributeError(name)
