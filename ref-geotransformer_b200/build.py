"""Builds lib/libgeotx_b200.so (the C-ABI CUDA library) in-tree with nvcc, for sm_100a only."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB_DIR = os.path.join(HERE, 'lib')
LIB_PATH = os.path.join(LIB_DIR, 'libgeotx_b200.so')
SOURCES = ['abi.cu', 'k0_probe.cu', 'k1_knn.cu', 'k2_gse_embed_fp32.cu', 'k2_gse_embed.cu', 'k3_rpe_scores.cu', 'k3_attention.cu', 'k3_attention_f16.cu',
           'k4_add_layernorm.cu', 'k5_superpoint_matching.cu', 'k6_point_to_node.cu',
           'k7_log_optimal_transport.cu', 'k8_weighted_procrustes.cu',
           'k9_radius_neighbors.cu', 'k10_grid_subsample.cu', 'k11_local_global_registration.cu', 'k12_patch_scores.cu', 'k2_gse_embed_table.cu', 'k2_gse_embed_cheb.cu', 'k4_post_attention.cu', 'k13_linear.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
              '-Xcompiler', '-fPIC']


def _nvcc():
    for cand in (shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found: the geotx_b200 CUDA library cannot be built')


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    os.makedirs(LIB_DIR, exist_ok=True)
    obj_dir = os.path.join(HERE, 'build')
    os.makedirs(obj_dir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cuh', '.h'))]
    headers.append(os.path.join(os.path.dirname(HERE), 'include', 'geotx_b200.h'))
    nvcc = _nvcc()
    objs = []
    procs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(obj_dir, src.replace('.cu', '.o'))
        objs.append(o)
        if force or _stale(o, [s] + headers):
            cmd = [nvcc] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + ['-c', s, '-o', o]
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out)
        if p.returncode != 0:
            raise RuntimeError(f'nvcc failed on {src}:\n{out}')
    if force or procs or _stale(LIB_PATH, objs):
        cmd = [nvcc, '-shared', '-o', LIB_PATH] + objs + ['-gencode', 'arch=compute_100a,code=sm_100a', '-lcudart']
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f'link failed:\n{r.stdout}')
    return LIB_PATH


if __name__ == '__main__':
    import sys
    print(build_library(force='--force' in sys.argv, verbose='-v' in sys.argv))
