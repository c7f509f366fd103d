"""matcha_b200 -- B200-native (sm_100a) drop-ins for the two data-parallel hot paths of
Matcha-TTS: monotonic alignment search and the CFM decoder solve.

Host code is Python/PyTorch (as in the reference); all arithmetic of the hot paths runs in
hand-written CUDA inside ``lib/libmatcha_b200.so``, reached through the C ABI declared in
``include/matcha_b200.h``.  There is no CPU fallback: importing a path without the built
library, or calling it without a CUDA device, raises.
"""
from . import alignment, decoder, flow_matching, monotonic_align, text_encoder  # noqa: F401
from . import matcha_tts  # noqa: F401
from ._lib import LibraryNotBuilt, MatchaB200Error, lib_path, load_library  # noqa: F401

__version__ = "0.1.0"
