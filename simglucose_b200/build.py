"""Compile the CUDA sources in csrc/ into simglucose_b200/lib/libsimglucose_b200.so (sm_100a only).

    python -m simglucose_b200.build [--force]

The kernels are compiled twice from the same source: once with -fmad=false (namespace sg_strict,
the default arithmetic: no FMA contraction, same accept/reject decisions as the scipy path) and once
with -fmad=true (namespace sg_fma, selected by SG_FLAG_FMA).  The library is built in-tree so it
travels with the repository snapshot; nothing is installed into site-packages.
"""
import os
import shutil
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, 'csrc')
LIB_DIR = os.path.join(_HERE, 'lib')
# A/B builds (tools/ab_variants.sh): SG_LIB_PATH names another output library, SG_NVCC_EXTRA adds compile flags
# (e.g. -DSG_EXP_TABLE=0); the product always loads the default path.
LIB_PATH = os.environ.get('SG_LIB_PATH') or os.path.join(LIB_DIR, 'libsimglucose_b200.so')
OBJ_DIR = os.path.join(LIB_DIR, 'obj' if not os.environ.get('SG_LIB_PATH') else
                       'obj_' + os.path.splitext(os.path.basename(LIB_PATH))[0])
EXTRA = os.environ.get('SG_NVCC_EXTRA', '').split()

ARCH = ['-gencode', 'arch=compute_100a,code=sm_100a']
COMMON = ['-O3', '-std=c++17', '-lineinfo', '-Xcompiler', '-fPIC', '-Xptxas', '-v']

UNITS = [
    # (object name, source, extra flags)
    ('sg_kernels_strict.o', 'sg_kernels.cu', ['-fmad=false', '-DSG_NS=sg_strict']),
    ('sg_kernels_fma.o', 'sg_kernels.cu', ['-fmad=true', '-DSG_NS=sg_fma', '-DSG_WITH_FP32']),
    ('sg_generators.o', 'sg_generators.cu', ['-fmad=true']),
    ('sg_coherence.o', 'sg_coherence.cu', ['-fmad=false']),
    ('sg_report.o', 'sg_report.cu', ['-fmad=false']),
    ('sg_noalloc.o', 'sg_noalloc.cu', ['-fmad=false']),
    ('sg_capi.o', 'sg_capi.cu', ['-fmad=false']),
]


def _nvcc():
    exe = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(exe):
        raise RuntimeError('nvcc not found: the CUDA extension cannot be built')
    return exe


def _sources():
    deps = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))]
    deps.append(os.path.join(_HERE, '..', 'include', 'simglucose_b200.h'))
    deps.append(os.path.abspath(__file__))
    return deps


def is_stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(d) > t for d in _sources() if os.path.exists(d))


def build(force=False, verbose=False):
    """Build the shared library if missing or older than its sources.  Returns its path."""
    if not force and not is_stale():
        return LIB_PATH
    nvcc = _nvcc()
    os.makedirs(OBJ_DIR, exist_ok=True)
    objs, log = [], []
    procs = []
    for obj, src, extra in UNITS:
        srcp = os.path.join(CSRC, src)
        if not os.path.exists(srcp):
            continue
        out = os.path.join(OBJ_DIR, obj)
        cmd = [nvcc] + ARCH + COMMON + extra + EXTRA + ['-c', srcp, '-o', out]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(out)
    for cmd, p in procs:
        out, _ = p.communicate()
        log.append(' '.join(cmd) + '\n' + out)
        if p.returncode != 0:
            raise RuntimeError('nvcc failed:\n' + log[-1])
    cmd = [nvcc] + ARCH + ['-shared', '-o', LIB_PATH] + objs
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log.append(' '.join(cmd) + '\n' + r.stdout)
    if r.returncode != 0:
        raise RuntimeError('link failed:\n' + log[-1])
    with open(os.path.join(LIB_DIR, 'build.log' if not os.environ.get('SG_LIB_PATH') else os.path.basename(LIB_PATH) + '.log'), 'w') as f:
        f.write('\n'.join(log))
    if verbose:
        print('\n'.join(log))
    return LIB_PATH


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
