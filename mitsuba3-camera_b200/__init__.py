"""lensrt — B200-native realistic-lens path tracer behind the Mitsuba 3 plugin API (the render hot path of
gitmon/mitsuba3-camera). Import as `import lensrt as mi` (repo-root shim) — the directory name carries a
hyphen, so it is loaded by path."""
from . import _capi as capi           # fails loudly if liblensrt.so has not been built
from . import lens_designs
from .film import ReconstructionFilter, ImageBlock, FilmDesc, develop_rgb
from .sensors import Film, Sensor, RealisticLensCamera, PerspectiveCamera

__version__ = "0.1.0"
