"""mitsuba3_nasa_b200 -- B200-native volumetric path tracing behind the mitsuba3-nasa plugin API.

    import mitsuba3_nasa_b200 as mi
    mi.set_variant('llvm_ad_mono_polarized')
    scene = mi.load_dict({...})
    stokes = mi.render_1(scene, spp=1 << 20)      # Mueller matrix, Stokes vector in column 0
"""
from .transforms import ScalarTransform4f, Transform4f  # noqa: F401
from .api import (set_variant, variant, variants, load_dict, load_string, load_file, traverse, render, render_1,  # noqa: F401
                  Scene, SceneParameters, Integrator, sample_tea_32)
from .scene import PluginError  # noqa: F401
from .exr import write_exr, read_exr  # noqa: F401
from .xmlscene import xml_to_dict, dict_to_xml, XMLError  # noqa: F401
from .ztest import z_test, images_agree  # noqa: F401
from . import dr  # noqa: F401,E402  (dr.enable_grad / dr.backward / dr.grad over torch autograd)
