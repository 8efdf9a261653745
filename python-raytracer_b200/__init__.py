"""B200-native render path behind the reference's own surface:
`scene_parser.load_scene(json)` in, `Scene.render(subimage, tasks)` out."""
from . import _abi, geometry, helperclasses, scene, scene_parser  # noqa: F401
from ._abi import B200RTError  # noqa: F401
from .scene import Scene, to_png_array  # noqa: F401
from .scene_parser import load_scene  # noqa: F401
