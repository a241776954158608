"""Import shim: the package directory is `deep-go_b200/` (named after the reference repo), which is not a
valid Python identifier, so `import deepgo_b200` resolves here and this module turns itself into that package."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "deep-go_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
