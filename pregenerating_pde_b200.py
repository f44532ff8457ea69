"""Import shim: the package directory is named `pregenerating-pde_b200` (not a Python identifier);
`import pregenerating_pde_b200` loads it under this importable name."""
import importlib.util as _u, os as _os, sys as _sys

_dir = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), 'pregenerating-pde_b200')
_spec = _u.spec_from_file_location('pregenerating_pde_b200', _os.path.join(_dir, '__init__.py'),
                                   submodule_search_locations=[_dir])
_mod = _u.module_from_spec(_spec)
_sys.modules['pregenerating_pde_b200'] = _mod
_spec.loader.exec_module(_mod)
