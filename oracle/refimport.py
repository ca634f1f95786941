"""TEST INFRASTRUCTURE ONLY -- import the REAL Python reference (pyamg 3.0.2 fork) straight from
/root/reference, without copying any of its sources: a meta-path finder serves
``pyamg.*`` modules from /root/reference/pyamg, patching four source lines IN MEMORY
 and supplying the two build-generated modules; the native module is
oracle/_ref/_amg_core.so (compiled from the reference sources by oracle/build_ref.sh).

Source of the package: /root/reference in the build container, else the installed copy under baseline/_ref
(oracle/install_ref.py; git-ignored, travels to the GPU box).  Used by tests/golden/make_golden.py to generate
fixtures, by tests that are skipped where neither exists, and by bench.py --impl reference (the reference arm).
"""
import builtins
import importlib
import importlib.abc
import importlib.machinery
import importlib.util
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
_INSTALLED = os.path.join(os.path.dirname(_HERE), 'baseline', '_ref')      # oracle/install_ref.py (travels to the GPU box)


def _ref_root():
    env = os.environ.get('PYAMG_REF_ROOT')
    if env:
        return env
    if os.path.isdir('/root/reference/pyamg'):
        return '/root/reference'
    return _INSTALLED


REF_ROOT = _ref_root()


def _native():
    for p in (os.path.join(_HERE, '_ref', '_amg_core.so'), os.path.join(_INSTALLED, '_amg_core.so')):
        if os.path.exists(p):
            return p
    return os.path.join(_HERE, '_ref', '_amg_core.so')

_PATCHES = {
    'aggregation/aggregate.py': [
        ('from matching import *', 'from .matching import *'),
    ],
    'gallery/stencil.py': [('diag[s] = 0', 'diag[tuple(s)] = 0')],
}


def available():
    return os.path.isdir(os.path.join(REF_ROOT, 'pyamg')) and os.path.exists(_native())


class _Loader(importlib.abc.SourceLoader):
    def __init__(self, fullname, path):
        self.fullname, self.path = fullname, path

    def get_filename(self, fullname):
        return self.path

    def get_data(self, path):
        with open(path, 'rb') as f:
            data = f.read()
        rel = os.path.relpath(path, os.path.join(REF_ROOT, 'pyamg'))
        if rel in _PATCHES:
            src = data.decode('utf-8')
            for old, new in _PATCHES[rel]:
                src = src.replace(old, new)
            if rel == 'aggregation/aggregate.py':   # the one py2 print statement (:380-381)
                import re
                src = re.sub(r'print "Warning - no left[^"]*"', 'print("no left near-nullspace")',
                             src, flags=re.S)
            data = src.encode('utf-8')
        return data

    def path_stats(self, path):
        raise OSError  # never write / read .pyc next to the read-only reference

    def set_data(self, *a, **k):
        pass


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if fullname != 'pyamg' and not fullname.startswith('pyamg.'):
            return None
        if fullname in ('pyamg.__config__', 'pyamg.version'):
            return importlib.machinery.ModuleSpec(fullname, _Generated(fullname))
        if fullname == 'pyamg.amg_core._amg_core':
            return importlib.util.spec_from_file_location(fullname, _native())
        rel = fullname.split('.')[1:]
        base = os.path.join(REF_ROOT, 'pyamg', *rel)
        if os.path.isdir(base) and os.path.exists(os.path.join(base, '__init__.py')):
            f = os.path.join(base, '__init__.py')
            return importlib.util.spec_from_file_location(
                fullname, f, loader=_Loader(fullname, f), submodule_search_locations=[base])
        if os.path.exists(base + '.py'):
            return importlib.util.spec_from_file_location(
                fullname, base + '.py', loader=_Loader(fullname, base + '.py'))
        return None


class _Generated(importlib.abc.Loader):
    def __init__(self, name):
        self.name = name

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        if self.name.endswith('__config__'):
            module.show = lambda: None
        else:
            module.version = module.short_version = module.full_version = '3.0.2'
            module.git_revision = 'unknown'
            module.release = True


def _shims():
    """Names removed from numpy/scipy since 2016 that the reference still uses."""
    import numpy as np
    import numpy.testing
    import scipy
    import scipy.linalg
    import scipy.sparse
    import scipy.sparse.linalg
    builtins.basestring = str
    for n, v in (('float', float), ('int', int), ('complex', complex), ('longfloat', np.longdouble)):
        if not hasattr(np, n):
            setattr(np, n, v)
    if not hasattr(np, 'find_common_type'):
        np.find_common_type = lambda a, b: np.result_type(*a, *b)
    if not hasattr(np, 'deprecate'):
        np.deprecate = lambda f=None, **k: (f if f is not None else (lambda g: g))
    scipy.rand = np.random.rand
    scipy.random = np.random
    for n in ('arange', 'ones', 'zeros', 'array', 'mod', 'eye', 'isscalar', 'hstack', 'ravel', 'sqrt',
              'conjugate', 'dot', 'diag', 'finfo', 'asarray', 'isreal', 'iscomplexobj', 'abs', 'real',
              'imag', 'cumsum', 'unique', 'empty', 'vstack', 'where', 'concatenate', 'multiply',
              'log10', 'ceil', 'floor', 'array_split', 'repeat', 'setdiff1d', 'argsort', 'sort',
              'zeros_like', 'ones_like', 'tile', 'kron', 'mat', 'asmatrix', 'rank', 'inner', 'sign',
              'amax', 'amin', 'pi', 'cos', 'sin', 'exp', 'log', 'sum', 'prod', 'mean', 'float64',
              'int32', 'complex128', 'linspace', 'meshgrid', 'ndarray', 'shape', 'size', 'reshape',
              'transpose', 'outer', 'identity', 'nonzero', 'maximum', 'minimum', 'isnan', 'isinf'):
        if not hasattr(scipy, n) and hasattr(np, n):
            setattr(scipy, n, getattr(np, n))
    if not hasattr(scipy, 'mat'):
        scipy.mat = np.asmatrix

    class _Tester:
        def test(self, *a, **k):
            pass
        bench = test
    if not hasattr(numpy.testing, 'Tester'):
        numpy.testing.Tester = _Tester
    try:
        import scipy.sparse.sputils as su
    except Exception:
        su = types.ModuleType('scipy.sparse.sputils')
        sys.modules['scipy.sparse.sputils'] = su
        scipy.sparse.sputils = su
    from scipy.sparse import _sputils
    for n in ('upcast', 'isscalarlike', 'isintlike', 'isshape', 'issequence', 'isdense', 'get_index_dtype'):
        if hasattr(_sputils, n) and not hasattr(su, n):
            setattr(su, n, getattr(_sputils, n))
    if not hasattr(scipy.linalg, 'calc_lwork'):
        m = types.ModuleType('scipy.linalg.calc_lwork')
        scipy.linalg.calc_lwork = m
        sys.modules['scipy.linalg.calc_lwork'] = m
    if not hasattr(scipy.linalg, 'pinv2'):
        scipy.linalg.pinv2 = scipy.linalg.pinv
    imp = types.ModuleType('imp')

    def _fm(*a, **k):
        raise ImportError
    imp.find_module = _fm
    sys.modules.setdefault('imp', imp)
    try:
        _is = importlib.import_module('scipy.sparse.linalg._isolve')
        _iu = importlib.import_module('scipy.sparse.linalg._isolve.utils')
        sys.modules['scipy.sparse.linalg.isolve'] = _is
        sys.modules['scipy.sparse.linalg.isolve.utils'] = _iu
        scipy.sparse.linalg.isolve = _is
    except Exception:
        pass
    for n in ('matplotlib', 'matplotlib.pyplot'):
        if n not in sys.modules:
            try:
                importlib.import_module(n)
            except Exception:
                sys.modules[n] = types.ModuleType(n)
    if not hasattr(sys.modules['matplotlib'], 'pyplot'):
        sys.modules['matplotlib'].pyplot = sys.modules['matplotlib.pyplot']
    from scipy.sparse._base import _spbase
    if not hasattr(_spbase, 'H'):
        _spbase.H = property(lambda self: self.conjugate().transpose())
    if not hasattr(_spbase, 'getH'):
        _spbase.getH = lambda self: self.conjugate().transpose()
    if not hasattr(_spbase, 'A'):
        _spbase.A = property(lambda self: self.toarray())


_installed = False


def import_pyamg():
    """Return the reference ``pyamg`` package (imported from /root/reference)."""
    global _installed
    if not available():
        raise ImportError('reference not available (no %s or no oracle/_ref/_amg_core.so)' % REF_ROOT)
    if not _installed:
        _shims()
        sys.meta_path.insert(0, _Finder())
        # amg_core.py falls back to a plain `import _amg_core`
        spec = importlib.util.spec_from_file_location('_amg_core', _native())
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        sys.modules['_amg_core'] = mod
        _installed = True
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        return importlib.import_module('pyamg')
