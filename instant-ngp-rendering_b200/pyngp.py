"""`import pyngp as ngp` for scripts written against the reference's Python binding
(src/python_api.cu:263-733, scripts/run.py): the enums and the Testbed class under the reference's
names, over libngp_b200.so. Put this directory on sys.path the way the reference's scripts put
build/ there (scripts/common.py:20-27):

    sys.path.insert(0, "<repo>/instant-ngp-rendering_b200"); import pyngp as ngp

What the binding has and this path does not (GUI / VR, SDF / volume modes, meshes, camera paths)
raises with a message that says so; nothing falls back to the CPU.
"""
import enum
import importlib
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
if os.path.dirname(_HERE) not in sys.path:
    sys.path.insert(0, os.path.dirname(_HERE))
_pkg = importlib.import_module(os.path.basename(_HERE))
_tb = importlib.import_module(os.path.basename(_HERE) + ".testbed")

NgpbError = _pkg.NgpbError


# python_api.cu:268-352; values are the reference's enumerators (common.h:79-160)
TestbedMode = enum.IntEnum("TestbedMode", {"Nerf": 0, "Sdf": 1, "Image": 2, "Volume": 3, "None": 5})
GroundTruthRenderMode = enum.IntEnum("GroundTruthRenderMode", {"Shade": 0, "Depth": 1})
RenderMode = enum.IntEnum("RenderMode", {"AO": 0, "Shade": 1, "Normals": 2, "Positions": 3, "Depth": 4, "Distortion": 5, "Cost": 6, "Slice": 7})
LossType = enum.IntEnum("LossType", {"L2": 0, "L1": 1, "Mape": 2, "Smape": 3, "Huber": 4, "LogL1": 5, "RelativeL2": 6})
LossType.SmoothL1 = LossType.Huber
NerfActivation = enum.IntEnum("NerfActivation", {"None": 0, "ReLU": 1, "Logistic": 2, "Exponential": 3})
ColorSpace = enum.IntEnum("ColorSpace", {"Linear": 0, "SRGB": 1})
TonemapCurve = enum.IntEnum("TonemapCurve", {"Identity": 0, "ACES": 1, "Hable": 2, "Reinhard": 3})
LensMode = enum.IntEnum("LensMode", {"Perspective": 0, "OpenCV": 1, "FTheta": 2, "LatLong": 3, "OpenCVFisheye": 4, "Equirectangular": 5})


class Testbed(_tb.Testbed):
    """pyngp.Testbed (python_api.cu:383-567). `mode` is read-only in the reference and follows the
    data that was loaded; the two-argument constructors load data and a network config right away."""

    def __init__(self, mode=None, data_path=None, network_config=None):
        super().__init__()
        if mode is not None and int(mode) not in (_tb.MODE_NERF, _tb.MODE_IMAGE, _tb.MODE_NONE):
            raise NgpbError("ngp-b200: only the NeRF and image modes are built (DESIGN.md section 7)")
        if data_path is not None:
            self.load_training_data(data_path)
        if network_config is not None:
            if isinstance(network_config, dict):
                self.reload_network_from_json(network_config)
            else:
                self.reload_network_from_file(network_config)

    @property
    def mode(self):
        return TestbedMode(self._mode)

    @mode.setter
    def mode(self, v):
        self._mode = int(v)

    @property
    def tonemap_curve(self):
        return TonemapCurve(int(self._get("tonemap_curve")))

    @tonemap_curve.setter
    def tonemap_curve(self, v):
        self._set("tonemap_curve", int(v))

    @property
    def color_space(self):
        return ColorSpace(_tb.Testbed.color_space.fget(self))

    @color_space.setter
    def color_space(self, v):
        _tb.Testbed.color_space.fset(self, int(v))
