"""Import shim: the package sources live in ``tensorhive-fixed_b200/`` (the name the build contract asks
for, which is not a valid Python identifier).  ``import tensorhive_fixed_b200`` resolves here and loads
the real package from that directory."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), 'tensorhive-fixed_b200')
__path__.insert(0, _real)
with open(_os.path.join(_real, '__init__.py')) as _f:
    exec(compile(_f.read(), _os.path.join(_real, '__init__.py'), 'exec'))
del _f
