"""TEST / BENCH INFRASTRUCTURE ONLY.  The "offline install" of the reference for bench.py's --impl reference arm.

The reference cannot be pip-installed (its setup.py needs numpy.distutils, gone from numpy 2; three fork files hold
Python-2 syntax: SURVEY 8c-3), so this does by hand what `pip install --target baseline/_ref /root/reference` would have
done for a pure-Python tree: the package's .py files (not dev/, not the test directories, not example data) are placed,
UNMODIFIED, under baseline/_ref/pyamg/ together with the compiled extension oracle/_ref/_amg_core.so.  baseline/_ref/ is
git-ignored (never part of this repository's history) and not gpurun-ignored, so the installed reference travels to the
GPU box, where /root/reference does not exist.  oracle/refimport.py imports it from there, applying its four
compatibility patches in memory exactly as it does for /root/reference.

    python oracle/install_ref.py          # build container only; a no-op elsewhere
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.environ.get('PYAMG_REF_SRC', '/root/reference')
DST = os.path.join(ROOT, 'baseline', '_ref')
SKIP_DIRS = {'dev', 'tests', 'example_data', '__pycache__'}


def install():
    pkg = os.path.join(SRC, 'pyamg')
    if not os.path.isdir(pkg):
        sys.stderr.write('install_ref: %s absent; keeping what is under %s\n' % (pkg, DST))
        return False
    out = os.path.join(DST, 'pyamg')
    if os.path.isdir(out):
        shutil.rmtree(out)
    n = 0
    for base, dirs, files in os.walk(pkg):
        dirs[:] = [d for d in dirs if d not in SKIP_DIRS]
        rel = os.path.relpath(base, pkg)
        os.makedirs(os.path.join(out, rel), exist_ok=True)
        for f in files:
            if f.endswith('.py'):
                shutil.copyfile(os.path.join(base, f), os.path.join(out, rel, f))
                n += 1
    so = os.path.join(HERE, '_ref', '_amg_core.so')
    if os.path.exists(so):
        shutil.copyfile(so, os.path.join(DST, '_amg_core.so'))
    with open(os.path.join(DST, 'INSTALLED_FROM'), 'w') as f:
        f.write('%s (pyamg 3.0.2 fork), %d .py files, unmodified; extension built by oracle/build_ref.sh\n' % (SRC, n))
    return True


if __name__ == '__main__':
    print('installed' if install() else 'not installed')
