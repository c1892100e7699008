"""Import alias for the hyphen-named package directory.

The product package lives in
``interactive-stylization-of-topographic-maps-using-prompt-driven-diffusion-models_b200/``
(the name the build contract prescribes; hyphens make it un-importable with a plain
``import`` statement).  This shim makes the same directory importable as ``topo_b200``:
sub-modules are resolved from the real directory through ``__path__`` so every module
has exactly one identity (``topo_b200.<name>``).
"""
import os as _os

REAL_PACKAGE_DIR = _os.path.join(
    _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
    "interactive-stylization-of-topographic-maps-using-prompt-driven-diffusion-models_b200",
)
__path__ = [REAL_PACKAGE_DIR]
_init = _os.path.join(REAL_PACKAGE_DIR, "__init__.py")
with open(_init) as _f:
    exec(compile(_f.read(), _init, "exec"))
