from .primitives import SDFSphere, SDFBox, SDFPlane, SDFLine, SDFDisk, SDFTorus
from .transformations import (SDFAffineTransformation, SDFSmoothUnion, SDFUnion,
                              SDFRounding, SDFOnion)
from .scene_registry import make_test_scene, make_test_scene2
from .synthetic import make_synthetic_scene
