"""ctypes binding of oracle/libbre_oracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  It wraps the plain-C CPU restatement in bre_oracle.c.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbre_oracle.so")


class Rng(C.Structure):
    _fields_ = [("state", C.c_uint64), ("inc", C.c_uint64)]


class Medium(C.Structure):
    _fields_ = [("kind", C.c_int), ("sigma_a", C.c_float * 3), ("sigma_s", C.c_float * 3),
                ("sigma_t", C.c_float * 3), ("g", C.c_float), ("nx", C.c_int), ("ny", C.c_int),
                ("nz", C.c_int), ("density", C.POINTER(C.c_float)), ("invMaxDensity", C.c_float),
                ("w2m", C.c_float * 16), ("tau_steps", C.c_int)]


class Photons(C.Structure):
    _fields_ = [("n", C.c_uint32), ("pos", C.POINTER(C.c_float)), ("power", C.POINTER(C.c_float)),
                ("dir", C.POINTER(C.c_float)), ("radius", C.POINTER(C.c_float))]


class Node(C.Structure):
    _fields_ = [("lo", C.c_float * 3), ("hi", C.c_float * 3), ("left", C.c_int32),
                ("right", C.c_int32), ("first", C.c_int32), ("count", C.c_int32), ("axis", C.c_int32)]


class Lbvh(C.Structure):
    _fields_ = [("n", C.c_uint32), ("order", C.POINTER(C.c_uint32)), ("codes", C.POINTER(C.c_uint32)),
                ("nodes", C.POINTER(Node)), ("leaf_lo", C.POINTER(C.c_float)),
                ("leaf_hi", C.POINTER(C.c_float)), ("cb_lo", C.c_float * 3), ("cb_hi", C.c_float * 3),
                ("abs_max", C.c_float)]


class CountersS(C.Structure):
    _fields_ = [("nodes_visited", C.c_uint64), ("disk_tests", C.c_uint64), ("hits", C.c_uint64)]


class LightS(C.Structure):
    _fields_ = [("kind", C.c_int), ("pos", C.c_float * 3), ("I", C.c_float * 3),
                ("axis", C.c_float * 3), ("cosTotalWidth", C.c_float), ("cosFalloffStart", C.c_float)]


class MeshS(C.Structure):
    _fields_ = [("n_tri", C.c_uint32), ("P", C.POINTER(C.c_float)), ("idx", C.POINTER(C.c_uint32)),
                ("Kd", C.POINTER(C.c_float)), ("attr", C.POINTER(C.c_uint8))]


def build_library():
    subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)


def load():
    if not os.path.exists(LIB_PATH):
        build_library()
    lib = C.CDLL(LIB_PATH)
    lib.orc_rng_float.restype = C.c_float
    lib.orc_rng_u32.restype = C.c_uint32
    lib.orc_phase_hg.restype = C.c_float
    lib.orc_phase_hg.argtypes = [C.c_float, C.c_float]
    lib.orc_hg_sample_p.restype = C.c_float
    lib.orc_grid_density.restype = C.c_float
    lib.orc_morton30.restype = C.c_uint32
    lib.orc_lbvh_build.restype = C.POINTER(Lbvh)
    lib.orc_trace_photons.restype = C.c_uint64
    lib.orc_trace_photons_mesh.restype = C.c_uint64
    return lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class Oracle:
    def __init__(self):
        self.lib = load()
        self.medium = Medium()
        self._keep = []
        self.tree = None
        self.ph = None

    # ---- media
    def set_medium_homogeneous(self, sa, ss, g):
        sa, ss = _f32(sa), _f32(ss)
        self.lib.orc_medium_homogeneous(C.byref(self.medium), _fp(sa), _fp(ss), C.c_float(g))

    def set_medium_grid(self, sa, ss, g, density, w2m):
        sa, ss, d, w = _f32(sa), _f32(ss), _f32(density), _f32(w2m).reshape(16)
        nz, ny, nx = d.shape
        self._keep += [d]
        self.lib.orc_medium_grid(C.byref(self.medium), _fp(sa), _fp(ss), C.c_float(g), nx, ny, nz, _fp(d), _fp(w))

    def set_tau_steps(self, steps):
        """Steps of the per-ray optical-depth table (default: orc_tau_steps_auto of the grid, like libbre.so)."""
        self.medium.tau_steps = int(steps)

    # ---- photons / tree
    def set_photons(self, pos, power, dirs, radius):
        pos, power, dirs = _f32(pos).reshape(-1, 3), _f32(power).reshape(-1, 3), _f32(dirs).reshape(-1, 3)
        n = pos.shape[0]
        rad = _f32(np.broadcast_to(np.asarray(radius, np.float32), (n,)).copy())
        self._arrays = (pos, power, dirs, rad)
        self.ph = Photons(n, _fp(pos), _fp(power), _fp(dirs), _fp(rad))
        self.free_tree()

    def free_tree(self):
        if self.tree is not None:
            self.lib.orc_lbvh_free(self.tree)
            self.tree = None

    def build(self):
        self.free_tree()
        self.tree = self.lib.orc_lbvh_build(C.byref(self.ph))
        return self.tree

    def tree_arrays(self):
        t = self.tree.contents
        n = t.n
        order = np.ctypeslib.as_array(t.order, (n,)).copy()
        codes = np.ctypeslib.as_array(t.codes, (n,)).copy()
        m = max(n - 1, 0)
        lo = np.zeros((m, 3), np.float32); hi = np.zeros((m, 3), np.float32)
        ch = np.zeros((m, 2), np.int32)
        if m:
            raw = np.ctypeslib.as_array(C.cast(t.nodes, C.POINTER(C.c_int32)), (m, C.sizeof(Node) // 4))
            lo[:] = raw[:, 0:3].view(np.float32); hi[:] = raw[:, 3:6].view(np.float32)
            ch[:] = raw[:, 6:8]
        return order, codes, lo, hi, ch

    def query(self, o_tmax, dirs, n_paths_norm, want_hits=False, brute=False, threads=0):
        o, d = _f32(o_tmax).reshape(-1, 4), _f32(dirs).reshape(-1, 3)
        n = o.shape[0]
        out = np.zeros((n, 3), np.float32)
        cnt = CountersS()
        offs = np.zeros(n + 1, np.uint64)
        po = offs.ctypes.data_as(C.POINTER(C.c_uint64)) if want_hits else None

        def call(idx_ptr, cap):
            if brute:
                return self.lib.orc_query_batch_brute(C.byref(self.ph), C.byref(self.medium), n, _fp(o), _fp(d),
                                                      C.c_float(n_paths_norm), _fp(out), po, idx_ptr,
                                                      C.c_uint64(cap), threads)
            return self.lib.orc_query_batch(self.tree, C.byref(self.ph), C.byref(self.medium), n, _fp(o), _fp(d),
                                            C.c_float(n_paths_norm), _fp(out), po, idx_ptr, C.c_uint64(cap),
                                            C.byref(cnt), threads)
        if not want_hits:
            call(None, 0)
            return out, dict(nodes_visited=cnt.nodes_visited, disk_tests=cnt.disk_tests, hits=cnt.hits)
        call(None, 1 << 62)
        total = int(offs[n])
        idx = np.zeros(max(total, 1), np.uint32)
        cnt = CountersS()
        call(idx.ctypes.data_as(C.POINTER(C.c_uint32)), total)
        return out, offs, idx[:total], dict(nodes_visited=cnt.nodes_visited, disk_tests=cnt.disk_tests, hits=cnt.hits)

    # ---- film
    def film_resolve(self, ld, it, scale=1.0):
        ld = _f32(ld).reshape(-1, 3)
        out = np.zeros_like(ld)
        self.lib.orc_film_resolve(_fp(ld), ld.shape[0], it, C.c_float(scale), _fp(out))
        return out

    # ---- photon tracing
    def trace(self, lights, blo, bhi, first, n_paths, max_depth, cap=None):
        arr = (LightS * len(lights))()
        for i, L in enumerate(lights):
            arr[i].kind = L.get("kind", 0)
            arr[i].pos = (C.c_float * 3)(*L["pos"]); arr[i].I = (C.c_float * 3)(*L["I"])
            arr[i].axis = (C.c_float * 3)(*L.get("axis", (0, 0, 1)))
            arr[i].cosTotalWidth = L.get("cos_total_width", 0.0)
            arr[i].cosFalloffStart = L.get("cos_falloff_start", 0.0)
        cap = cap or n_paths * max_depth
        pos = np.zeros((cap, 3), np.float32); pw = np.zeros((cap, 3), np.float32)
        dr = np.zeros((cap, 3), np.float32); pid = np.zeros(cap, np.uint32)
        blo, bhi = _f32(blo), _f32(bhi)
        n = self.lib.orc_trace_photons(C.byref(self.medium), arr, len(lights), _fp(blo), _fp(bhi),
                                       C.c_uint64(first), C.c_uint64(n_paths), max_depth, _fp(pos), _fp(pw),
                                       _fp(dr), pid.ctypes.data_as(C.POINTER(C.c_uint32)), C.c_uint64(cap))
        n = min(int(n), cap)
        return pos[:n], pw[:n], dr[:n], pid[:n]

    def _lights(self, lights):
        arr = (LightS * len(lights))()
        for i, L in enumerate(lights):
            arr[i].kind = L.get("kind", 0)
            arr[i].pos = (C.c_float * 3)(*L["pos"]); arr[i].I = (C.c_float * 3)(*L["I"])
            arr[i].axis = (C.c_float * 3)(*L.get("axis", (0, 0, 1)))
            arr[i].cosTotalWidth = L.get("cos_total_width", 0.0)
            arr[i].cosFalloffStart = L.get("cos_falloff_start", 0.0)
        return arr

    # ---- surfaces
    def set_mesh(self, P, idx, Kd, attr=None):
        P, Kd = _f32(P).reshape(-1, 3), _f32(Kd).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, np.uint32).reshape(-1, 3)
        a = np.zeros(idx.shape[0], np.uint8) if attr is None else np.ascontiguousarray(attr, np.uint8)
        self._mesh_arrays = (P, idx, Kd, a)
        self.mesh = MeshS(idx.shape[0], _fp(P), idx.ctypes.data_as(C.POINTER(C.c_uint32)), _fp(Kd),
                          a.ctypes.data_as(C.POINTER(C.c_uint8)))

    def trace_mesh(self, lights, first, n_paths, max_depth, lights_in_medium=True, cap=None):
        arr = self._lights(lights)
        cap = cap or n_paths * max_depth
        pos = np.zeros((cap, 3), np.float32); pw = np.zeros((cap, 3), np.float32)
        dr = np.zeros((cap, 3), np.float32); pid = np.zeros(cap, np.uint32)
        n = self.lib.orc_trace_photons_mesh(C.byref(self.medium), arr, len(lights), C.byref(self.mesh),
                                            int(lights_in_medium), C.c_uint64(first), C.c_uint64(n_paths),
                                            max_depth, _fp(pos), _fp(pw), _fp(dr),
                                            pid.ctypes.data_as(C.POINTER(C.c_uint32)), C.c_uint64(cap))
        n = min(int(n), cap)
        return pos[:n], pw[:n], dr[:n], pid[:n]

    def halton_camera_samples(self, x0, y0, x1, y1, sample_num):
        """(dimension 0, dimension 1) of the reference's HaltonSampler per pixel of [x0,x1) x [y0,y1), row major."""
        out = np.zeros(((x1 - x0) * (y1 - y0), 2), np.float32)
        self.lib.orc_halton_camera_samples(int(x0), int(y0), int(x1), int(y1), C.c_int64(int(sample_num)), _fp(out))
        return out

    def camera_rays_px(self, r2c, c2w, x0, y0, w, h, offs_xy, use_mesh=True, camera_in_medium=True):
        r2c, c2w, offs = _f32(r2c).reshape(16), _f32(c2w).reshape(16), _f32(offs_xy).reshape(-1)
        o, d = np.zeros((w * h, 4), np.float32), np.zeros((w * h, 3), np.float32)
        mesh = C.byref(self.mesh) if use_mesh else None
        self.lib.orc_camera_rays_px(_fp(r2c), _fp(c2w), x0, y0, w, h, C.c_float(0), C.c_float(0), _fp(offs), mesh,
                                    int(camera_in_medium), _fp(o), _fp(d))
        return o, d

    def camera_rays(self, r2c, c2w, x0, y0, w, h, offx, offy, use_mesh=True, camera_in_medium=True):
        r2c, c2w = _f32(r2c).reshape(16), _f32(c2w).reshape(16)
        o = np.zeros((w * h, 4), np.float32); d = np.zeros((w * h, 3), np.float32)
        mesh = C.byref(self.mesh) if use_mesh else None
        self.lib.orc_camera_rays(_fp(r2c), _fp(c2w), x0, y0, w, h, C.c_float(offx), C.c_float(offy), mesh,
                                 int(camera_in_medium), _fp(o), _fp(d))
        return o, d

    # ---- synthetic workloads
    def synth_photons_uniform(self, n):
        pos = np.zeros((n, 3), np.float32); pw = np.zeros((n, 3), np.float32); dr = np.zeros((n, 3), np.float32)
        self.lib.orc_synth_photons_uniform(n, _fp(pos), _fp(pw), _fp(dr))
        return pos, pw, dr

    def synth_rays_pinhole(self, w, h):
        o = np.zeros((w * h, 4), np.float32); d = np.zeros((w * h, 3), np.float32)
        self.lib.orc_synth_rays_pinhole(w, h, _fp(o), _fp(d))
        return o, d

    def synth_rays_incoherent(self, n):
        o = np.zeros((n, 4), np.float32); d = np.zeros((n, 3), np.float32)
        self.lib.orc_synth_rays_incoherent(n, _fp(o), _fp(d))
        return o, d
