"""lensrt — B200-native realistic-lens path tracer behind the Mitsuba 3 plugin API (the render hot path of
gitmon/mitsuba3-camera). Import as `import lensrt as mi` (repo-root shim) — the directory name carries a
hyphen, so it is loaded by path."""
from . import _capi as capi           # fails loudly if liblensrt.so has not been built
from . import lens_designs
from .film import ReconstructionFilter, ImageBlock, FilmDesc, develop_rgb
from .sensors import Film, Sensor, RealisticLensCamera, PerspectiveCamera

__version__ = "0.1.0"
from .transform import Transform4f, ScalarTransform4f
from .scene import (Scene, SceneData, PathIntegrator, PRBIntegrator, PTracerIntegrator, PtraceParams, IndependentSampler, load_dict, render, cornell_box, bsdf_row,
                    film_desc, RenderParams, SceneDesc, load_obj, load_ply, scene_data_from_dict, render_backward, traverse,
                    SceneParameters)
from . import scene, workloads, lens_design, parallel, imageio, ad, mesh_lens
from .xml import load_file, load_string, xml_to_dict, XMLError
from . import imageio as util          # mi.util.write_bitmap (src/python/python/util.py) for .exr / .pfm
from .lens_design import spot_render, spot_render_backward, spot_image, rms_loss, field_source

_variant = "b200_ad_rgb"


def variants():
    return ["b200_ad_rgb"]


def variant():
    return _variant


def set_variant(*names):
    """mi.set_variant: this build has one variant; reference names of the *_rgb family are accepted."""
    for n in names:
        if n in ("b200_ad_rgb", "cuda_ad_rgb", "llvm_ad_rgb", "scalar_rgb", "cuda_rgb", "llvm_rgb"):
            return
    raise ImportError(f"variant(s) {names} not available; this build provides 'b200_ad_rgb'")
