"""ol_dsp_b200: B200-native banks of ol_dsp units.  All compute is in libolbank.so (sm_100a CUDA)."""
from . import _lib  # noqa: F401
from ._lib import (KIND_BIQUAD, KIND_LPF, KIND_PORT, LAYOUT_FRAMES, LAYOUT_PLANAR, MODE_EXACT, MODE_FAST, MODE_FMA, OlbError)  # noqa: F401
from .chorus import ChorusBank, StereoChorusBank  # noqa: F401
from .sampler import SamplerBank  # noqa: F401
from .dattorro import DattorroBank  # noqa: F401
from .fx import FxBank  # noqa: F401
from .voice import Polyvoice, VoiceBank, VoiceMap  # noqa: F401
from .sharding import Comm, shard_range  # noqa: F401
from .biquad import BiquadBank, design  # noqa: F401
from .workloads import c2_coefs, noise_device  # noqa: F401
