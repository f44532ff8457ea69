"""Builds libcno_b200.so (sm_100a) in-tree with plain nvcc; no torch headers involved.

    python pregenerating-pde_b200/build.py [--force] [--quick]

Objects go to build/cno_b200/, the shared library next to this file.  Replaces the reference's JIT
plugin build (torch_utils/custom_ops.py:53-149): ahead of time, one library, no per-GPU cache.
"""
import os, subprocess, sys, concurrent.futures, shutil

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OBJ = os.path.join(os.path.dirname(HERE), 'build', 'cno_b200')
LIB = os.path.join(HERE, 'libcno_b200.so')
SOURCES = ['capi.cu', 'bias_act.cu', 'upfirdn2d.cu', 'flrelu_act.cu', 'cin.cu', 'adamw.cu', 'layout.cu', 'flrelu_fused.cu',
           'flrelu_fused_f32.cu', 'flrelu_fused_bf16.cu', 'flrelu_fused_f16.cu',
           'flrelu_plane_f32.cu', 'flrelu_plane_bf16.cu', 'flrelu_plane_f16.cu'] + \
          [f'flrelu_fused_{t}_p{i}.cu' for t in ('f32', 'bf16', 'f16') for i in range(5)]
NVCC_FLAGS = ['-std=c++17', '-O3', '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo',
              '-Xcompiler', '-fPIC', '--threads', '2']


def nvcc():
    exe = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'
    if not os.path.exists(exe):
        raise RuntimeError('nvcc not found')
    return exe


def newest_header():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.h', '.cuh'))]
    hs.append(os.path.join(os.path.dirname(HERE), 'include', 'cno_b200.h'))
    return max(os.path.getmtime(h) for h in hs)


def compile_one(src, force, extra):
    obj = os.path.join(OBJ, src.replace('.cu', '.o'))
    path = os.path.join(CSRC, src)
    if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(path), newest_header()):
        return obj, False
    subprocess.check_call([nvcc()] + NVCC_FLAGS + extra + ['-c', path, '-o', obj])
    return obj, True


def build(force=False, quick=False, verbose=True):
    os.makedirs(OBJ, exist_ok=True)
    extra = ['-DFLR_QUICK'] if quick else []
    sources = [s for s in SOURCES if not (quick and '_p' in s and s.startswith('flrelu_fused_'))]
    sources.sort(key=lambda s: 0 if '_p' in s else 1)          # the long compilations first
    stamp = os.path.join(OBJ, 'quick.stamp')
    if quick != os.path.exists(stamp):
        force = True
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as ex:
        results = list(ex.map(lambda s: compile_one(s, force, extra), sources))
    if quick:
        open(stamp, 'w').close()
    elif os.path.exists(stamp):
        os.remove(stamp)
    objs = [o for o, _ in results]
    if any(changed for _, changed in results) or not os.path.exists(LIB):
        subprocess.check_call([nvcc(), '-shared', '-o', LIB] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '--cudart', 'static'])
        if verbose:
            print(f'built {LIB} ({os.path.getsize(LIB) >> 10} KiB)')
    elif verbose:
        print(f'{LIB} is up to date')
    return LIB


if __name__ == '__main__':
    build(force='--force' in sys.argv, quick='--quick' in sys.argv)
