"""Import alias: the package directory is named ``vq-vae-hmm-model_b200`` (not a valid Python
identifier), so this module turns itself into that package: ``import vhmm_b200`` and
``import vhmm_b200.functional`` resolve into the hyphenated directory."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "vq-vae-hmm-model_b200")]
__package__ = "vhmm_b200"
if globals().get("__spec__") is not None:
    __spec__.submodule_search_locations = __path__
__file__ = _os.path.join(__path__[0], "__init__.py")
with open(__file__) as _f:
    exec(compile(_f.read(), __file__, "exec"))
del _f, _os
