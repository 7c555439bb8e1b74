"""B200-native drop-in for the FFT Gaussian-blur / FastBoxBlur / DFT_image hot path of
michelerenzullo/Blur_algorithms.  See DESIGN.md and include/blur_b200.h."""
from .api import *  # noqa: F401,F403
from .api import __all__  # noqa: F401
