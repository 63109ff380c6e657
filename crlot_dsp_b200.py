"""Import shim: the package directory is named `crlot-dsp_b200/` (the project's name), which
is not a valid Python identifier.  `import crlot_dsp_b200` resolves here and behaves as that
package (``__path__`` points into the directory, so ``crlot_dsp_b200.capi`` etc. import
normally)."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "crlot-dsp_b200")]
__package__ = __name__
if __spec__ is not None:
    __spec__.submodule_search_locations = __path__
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
