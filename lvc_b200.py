"""Import shim: ``import lvc_b200`` loads the package that lives in
``ai-laryngeal-video-based-classifier_b200/`` (the directory name the project layout
prescribes is not a valid Python identifier)."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)),
                          "ai-laryngeal-video-based-classifier_b200")]
__package__ = __name__
if __spec__ is not None:
    __spec__.submodule_search_locations = __path__
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
del _f
