"""mio_tts_cpp_b200 -- B200-native MioCodec decode + iSTFT behind the reference's mio-tts-lib C API.

The product is the C-ABI shared library ``libmio_tts_b200.so`` (sources in ``csrc/``, header in
``include/mio_tts_b200.h``).  This package is only its Python-side mirror: a ctypes binding whose method names
and error behaviour follow ``/root/reference/src/mio-tts-lib.h`` so tests read like calls into the reference.
Nothing here computes anything: without the CUDA library every call fails loudly.
"""
from .lib import MioTTS, MioTtsError, MioTtsParams, lib_path, load_library  # noqa: F401

__all__ = ["MioTTS", "MioTtsError", "MioTtsParams", "lib_path", "load_library"]
