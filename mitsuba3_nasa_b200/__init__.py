"""mitsuba3_nasa_b200 -- B200-native volumetric path tracing behind the mitsuba3-nasa plugin API."""
from .transforms import ScalarTransform4f, Transform4f  # noqa: F401
