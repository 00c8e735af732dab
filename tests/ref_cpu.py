"""ctypes access to oracle/_ref/libgeotx_ref_cpu.so: the REFERENCE's own CPU collate operators compiled from
/root/reference by oracle/build_ref.sh (test infrastructure; see oracle/ref_cpu_shim.cpp)."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, 'oracle', '_ref', 'libgeotx_ref_cpu.so')
_lib = None


def available():
    return os.path.exists(SO) or os.path.isdir('/root/reference/geotransformer/extensions')


def load():
    global _lib
    if _lib is None:
        if os.path.isdir('/root/reference/geotransformer/extensions'):
            srcs = [os.path.join(ROOT, 'oracle', f) for f in ('ref_cpu_shim.cpp', 'build_ref.sh')]
            if not os.path.exists(SO) or any(os.path.getmtime(s) > os.path.getmtime(SO) for s in srcs):
                subprocess.run(['bash', os.path.join(ROOT, 'oracle', 'build_ref.sh')], check=True, stdout=subprocess.DEVNULL)
        lib = ctypes.CDLL(SO)
        lib.geotx_ref_radius_neighbors.restype = ctypes.c_int64
        lib.geotx_ref_grid_subsampling.restype = ctypes.c_int64
        _lib = lib
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def radius_neighbors(q_points, s_points, q_lengths, s_lengths, radius):
    """numpy in, numpy out: the reference's ext.radius_neighbors (untruncated (Nq, max_count) int64)."""
    lib = load()
    q = np.ascontiguousarray(q_points, np.float32)
    s = np.ascontiguousarray(s_points, np.float32)
    ql, sl = np.asarray(q_lengths, np.int64), np.asarray(s_lengths, np.int64)
    width = lib.geotx_ref_radius_neighbors(_ptr(q), ctypes.c_int64(q.shape[0]), _ptr(s), ctypes.c_int64(s.shape[0]),
                                           _ptr(ql), _ptr(sl), ctypes.c_int(len(ql)), ctypes.c_float(radius))
    out = np.empty((q.shape[0], width), np.int64)
    lib.geotx_ref_radius_neighbors_copy(_ptr(out))
    return out


def grid_subsampling(points, lengths, voxel_size):
    """The reference's ext.grid_subsampling: (s_points (M,3) float32, s_lengths (B,) int64)."""
    lib = load()
    p = np.ascontiguousarray(points, np.float32)
    ln = np.asarray(lengths, np.int64)
    m = lib.geotx_ref_grid_subsampling(_ptr(p), ctypes.c_int64(p.shape[0]), _ptr(ln), ctypes.c_int(len(ln)),
                                       ctypes.c_float(voxel_size))
    sp = np.empty((m, 3), np.float32)
    sl = np.empty(len(ln), np.int64)
    lib.geotx_ref_grid_subsampling_copy(_ptr(sp), _ptr(sl))
    return sp, sl


def precompute_data_stack_mode(points, lengths, num_stages, voxel_size, radius, neighbor_limits):
    """geotransformer/utils/data.py:13-81 on the reference's own operators (numpy arrays; the truncation of
    modules/ops/radius_search.py:24-27 included)."""
    def search(q, s, ql, sl, r, limit):
        return radius_neighbors(q, s, ql, sl, r)[:, :limit]
    points_list, lengths_list, neighbors, subsampling, upsampling = [], [], [], [], []
    for i in range(num_stages):
        if i > 0:
            points, lengths = grid_subsampling(points, lengths, voxel_size)
        points_list.append(points)
        lengths_list.append(np.asarray(lengths, np.int64))
        voxel_size *= 2
    for i in range(num_stages):
        cp, cl = points_list[i], lengths_list[i]
        neighbors.append(search(cp, cp, cl, cl, radius, neighbor_limits[i]))
        if i < num_stages - 1:
            sp, sl = points_list[i + 1], lengths_list[i + 1]
            subsampling.append(search(sp, cp, sl, cl, radius, neighbor_limits[i]))
            upsampling.append(search(cp, sp, cl, sl, radius * 2, neighbor_limits[i + 1]))
        radius *= 2
    return {'points': points_list, 'lengths': lengths_list, 'neighbors': neighbors, 'subsampling': subsampling,
            'upsampling': upsampling}
