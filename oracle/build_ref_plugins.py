"""Builds the reference's OWN CUDA plugins (bias_act / upfirdn2d / filtered_lrelu) for sm_100a -- TEST / BASELINE
INFRASTRUCTURE, not product code.

    python oracle/build_ref_plugins.py [/root/reference]

The sources are compiled where they lie under <reference>/CNO_Experiments/torch_utils/ops (nothing is copied into this
repository); only the built extension modules land in oracle/_ref/plugins/ (git-ignored, but shipped to the GPU box by
gpurun like our own libcno_b200.so).  This is the recipe of the reference's loader (torch_utils/custom_ops.py:53-149:
torch.utils.cpp_extension.load with `--use_fast_math`, filtered_lrelu.py:15-25, upfirdn2d.py:16-26, bias_act.py:27-37)
with the architecture pinned to 10.0a, because the build container has no GPU for the JIT to query.

Users: tests/test_ref_cuda_plugin.py (second, GPU-side oracle: outputs and sign-mask bytes bit for bit) and
bench_sweep.py (`ref_cuda_us` column: the StyleGAN3 kernels recompiled for Blackwell).
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, '_ref', 'plugins')

PLUGINS = {
    'bias_act_plugin': ['bias_act.cpp', 'bias_act.cu'],
    'upfirdn2d_plugin': ['upfirdn2d.cpp', 'upfirdn2d.cu'],
    'filtered_lrelu_plugin': ['filtered_lrelu.cpp', 'filtered_lrelu_wr.cu', 'filtered_lrelu_rd.cu', 'filtered_lrelu_ns.cu'],
}


def build(ref_root='/root/reference', verbose=False):
    os.environ['TORCH_CUDA_ARCH_LIST'] = '10.0a'
    os.environ.setdefault('MAX_JOBS', str(os.cpu_count() or 4))
    import torch.utils.cpp_extension as ext
    src_dir = os.path.join(ref_root, 'CNO_Experiments', 'torch_utils', 'ops')
    built = []
    for name, files in PLUGINS.items():
        bdir = os.path.join(OUT, name)
        os.makedirs(bdir, exist_ok=True)
        so = os.path.join(bdir, name + '.so')
        if os.path.exists(so):
            built.append(so)
            continue
        ext.load(name=name, sources=[os.path.join(src_dir, f) for f in files], build_directory=bdir,
                 extra_cuda_cflags=['--use_fast_math', '--allow-unsupported-compiler'], verbose=verbose, is_python_module=False)
        built.append(so)
    return built


def load(name):
    """imports a built plugin from oracle/_ref/plugins (returns None when it was never built)"""
    import importlib.util
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    so = os.path.join(OUT, name, name + '.so')
    if not os.path.exists(so):
        return None
    spec = importlib.util.spec_from_file_location(name, so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == '__main__':
    root = sys.argv[1] if len(sys.argv) > 1 else '/root/reference'
    for so in build(root, verbose='-v' in sys.argv):
        print(so, os.path.getsize(so) >> 10, 'KiB')
