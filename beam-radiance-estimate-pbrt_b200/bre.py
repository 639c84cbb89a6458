"""ctypes binding of libbre.so (include/bre.h) -- the call a Python user makes.

This is plumbing only: every function forwards to the C ABI, which runs hand-written
sm_100a CUDA kernels.  There is no CPU fallback: `Bre()` raises if the library or a CUDA
device is missing.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BRE_LIB_PATH") or os.path.join(_HERE, "libbre.so")   # override: kernel-variant experiments

BRE_OK = 0
ERR_NAMES = {0: "BRE_OK", -1: "BRE_ERR_INVALID", -2: "BRE_ERR_NO_DEVICE", -3: "BRE_ERR_CUDA",
             -4: "BRE_ERR_NOMEM", -5: "BRE_ERR_UNSUPPORTED", -6: "BRE_ERR_NCCL"}

# every symbol include/bre.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "bre_abi_version", "bre_create", "bre_destroy", "bre_last_error", "bre_get_stream",
    "bre_synchronize", "bre_nccl_unique_id", "bre_set_medium_homogeneous", "bre_set_medium_grid",
    "bre_set_lights", "bre_set_bounds", "bre_set_triangles", "bre_trace_photons",
    "bre_upload_photons", "bre_clear_photons", "bre_photon_count", "bre_download_photons",
    "bre_set_radius", "bre_allgather_photons", "bre_build", "bre_download_tree", "bre_query",
    "bre_query_d", "bre_count", "bre_count_d", "bre_get_counters", "bre_film_create",
    "bre_film_clear", "bre_film_accumulate", "bre_film_accumulate_d", "bre_film_resolve", "bre_film_mean",
    "bre_query_to_film", "bre_set_timing", "bre_get_timings", "bre_shard_range", "bre_tile_owner",
    "bre_set_triangle_attrs", "bre_set_lights_in_medium", "bre_set_camera", "bre_camera_rays",
    "bre_query_camera_to_film", "bre_last_build_kind",
    "bre_trace_exchange_photons", "bre_alloc_pinned", "bre_free_pinned", "bre_stage_rays", "bre_query_staged_to_film", "bre_save_photons", "bre_load_photons", "bre_film_save", "bre_film_load",
    "bre_film_resolve_static", "bre_camera_rays_halton", "bre_stage_camera_halton", "bre_query_camera_halton_to_film", "bre_set_ray_grid", "bre_set_tile_sharding", "bre_set_photon_sharding", "bre_film_gather", "bre_stage_camera", "bre_stage_rays_d",
]


class BreError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{ERR_NAMES.get(code, code)}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("device", C.c_int), ("n_ranks", C.c_int), ("rank", C.c_int),
                ("nccl_unique_id", C.c_void_p), ("max_leaf_photons", C.c_int),
                ("tau_steps", C.c_int)]


class Light(C.Structure):
    _fields_ = [("kind", C.c_int), ("pos", C.c_float * 3), ("I", C.c_float * 3),
                ("axis", C.c_float * 3), ("cos_total_width", C.c_float),
                ("cos_falloff_start", C.c_float)]


class Counters(C.Structure):
    _fields_ = [("nodes_visited", C.c_uint64), ("disk_tests", C.c_uint64), ("hits", C.c_uint64),
                ("bytes_algorithmic", C.c_uint64), ("fast_nodes", C.c_uint64),
                ("fast_disk_tests", C.c_uint64), ("leaf_pairs", C.c_uint64),
                ("pair_queue_overflow", C.c_uint64)]


class Camera(C.Structure):
    _fields_ = [("raster_to_camera", C.c_float * 16), ("camera_to_world", C.c_float * 16),
                ("x0", C.c_int32), ("y0", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("in_medium", C.c_int32)]


class HitDump(C.Structure):
    _fields_ = [("offsets", C.POINTER(C.c_uint64)), ("indices", C.POINTER(C.c_uint32)),
                ("capacity", C.c_uint64), ("total", C.c_uint64)]


def load_library(path=LIB_PATH):
    if not os.path.exists(path):
        raise OSError(f"{path} is missing: build it with `make lib` (python __graft_entry__.py build)")
    lib = C.CDLL(path)
    lib.bre_last_error.restype = C.c_char_p
    lib.bre_get_stream.restype = C.c_void_p
    lib.bre_alloc_pinned.restype = C.c_void_p
    lib.bre_alloc_pinned.argtypes = [C.c_size_t]
    lib.bre_free_pinned.restype = None
    lib.bre_free_pinned.argtypes = [C.c_void_p]
    lib.bre_shard_range.restype = None
    return lib


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


class Bre:
    """One context = one GPU (one rank)."""

    def __init__(self, device=0, n_ranks=1, rank=0, nccl_unique_id=None, max_leaf_photons=0,
                 tau_steps=0, lib=None):
        self.lib = lib or load_library()
        self.ctx = C.c_void_p()
        self._uid = None
        cfg = Config(device, n_ranks, rank, None, max_leaf_photons, tau_steps)
        if nccl_unique_id is not None:
            self._uid = C.create_string_buffer(bytes(nccl_unique_id), 128)
            cfg.nccl_unique_id = C.cast(self._uid, C.c_void_p)
        s = self.lib.bre_create(C.byref(self.ctx), C.byref(cfg))
        if s != BRE_OK:
            self.ctx = C.c_void_p()
            raise BreError(s, "bre_create failed (libbre.so has no CPU fallback; a CUDA device is required)")

    def _ck(self, s):
        if s != BRE_OK:
            raise BreError(s, self.lib.bre_last_error(self.ctx).decode())

    def close(self):
        if self.ctx:
            self.lib.bre_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return self.lib.bre_get_stream(self.ctx)

    def synchronize(self):
        self._ck(self.lib.bre_synchronize(self.ctx))

    # ---- scene
    def set_medium_homogeneous(self, sigma_a, sigma_s, g):
        sa, ss = _f32(sigma_a, (3,)), _f32(sigma_s, (3,))
        self._ck(self.lib.bre_set_medium_homogeneous(self.ctx, _fp(sa), _fp(ss), C.c_float(g)))

    def set_medium_grid(self, sigma_a, sigma_s, g, density, world_to_medium):
        sa, ss = _f32(sigma_a, (3,)), _f32(sigma_s, (3,))
        d = _f32(density)
        nz, ny, nx = d.shape
        w = _f32(world_to_medium, (16,))
        self._ck(self.lib.bre_set_medium_grid(self.ctx, _fp(sa), _fp(ss), C.c_float(g), nx, ny, nz,
                                              _fp(d), _fp(w)))

    def set_lights(self, lights):
        arr = (Light * len(lights))()
        for i, L in enumerate(lights):
            arr[i].kind = L.get("kind", 0)
            arr[i].pos = (C.c_float * 3)(*L["pos"])
            arr[i].I = (C.c_float * 3)(*L["I"])
            arr[i].axis = (C.c_float * 3)(*L.get("axis", (0, 0, 1)))
            arr[i].cos_total_width = L.get("cos_total_width", 0.0)
            arr[i].cos_falloff_start = L.get("cos_falloff_start", 0.0)
        self._ck(self.lib.bre_set_lights(self.ctx, len(lights), arr))

    def set_bounds(self, lo, hi):
        lo, hi = _f32(lo, (3,)), _f32(hi, (3,))
        self._ck(self.lib.bre_set_bounds(self.ctx, _fp(lo), _fp(hi)))

    def set_triangles(self, P, idx, Kd, attrs=None):
        P, Kd = _f32(P), _f32(Kd)
        idx = np.ascontiguousarray(idx, dtype=np.uint32)
        n = idx.size // 3
        self._ck(self.lib.bre_set_triangles(self.ctx, n, _fp(P), idx.ctypes.data_as(C.POINTER(C.c_uint32)), _fp(Kd)))
        if attrs is not None and n:
            a = np.ascontiguousarray(attrs, dtype=np.uint8)
            self._ck(self.lib.bre_set_triangle_attrs(self.ctx, n, a.ctypes.data_as(C.POINTER(C.c_uint8))))

    def set_lights_in_medium(self, flag):
        self._ck(self.lib.bre_set_lights_in_medium(self.ctx, int(flag)))

    def set_camera(self, raster_to_camera, camera_to_world, x0, y0, width, height, in_medium=True):
        cam = Camera()
        cam.raster_to_camera = (C.c_float * 16)(*_f32(raster_to_camera, (16,)))
        cam.camera_to_world = (C.c_float * 16)(*_f32(camera_to_world, (16,)))
        cam.x0, cam.y0, cam.width, cam.height, cam.in_medium = x0, y0, width, height, int(in_medium)
        self._ck(self.lib.bre_set_camera(self.ctx, C.byref(cam)))
        self._cam_n = width * height

    def camera_rays(self, offx=0.5, offy=0.5):
        n = getattr(self, "_cam_n", 0)     # 0: no camera yet -> the library reports the error
        o = np.zeros((n, 4), np.float32); d = np.zeros((n, 3), np.float32)
        self._ck(self.lib.bre_camera_rays(self.ctx, C.c_float(offx), C.c_float(offy), _fp(o), _fp(d)))
        return o, d

    def camera_rays_halton(self, sample_num):
        n = getattr(self, "_cam_n", 0)
        o = np.zeros((n, 4), np.float32); d = np.zeros((n, 3), np.float32)
        self._ck(self.lib.bre_camera_rays_halton(self.ctx, C.c_int64(int(sample_num)), _fp(o), _fp(d)))
        return o, d

    def stage_camera_halton(self, sample_num):
        self._ck(self.lib.bre_stage_camera_halton(self.ctx, C.c_int64(int(sample_num))))

    def query_camera_halton_to_film(self, n_paths_norm, sample_num):
        self._ck(self.lib.bre_query_camera_halton_to_film(self.ctx, C.c_int64(int(sample_num)), C.c_float(n_paths_norm)))

    def stage_camera(self, offx=0.5, offy=0.5):
        self._ck(self.lib.bre_stage_camera(self.ctx, C.c_float(offx), C.c_float(offy)))

    def stage_rays_d(self, n, rays_o_ptr, rays_d_ptr):
        self._ck(self.lib.bre_stage_rays_d(self.ctx, n, C.c_void_p(rays_o_ptr), C.c_void_p(rays_d_ptr)))

    def query_camera_to_film(self, n_paths_norm, offx=0.5, offy=0.5):
        self._ck(self.lib.bre_query_camera_to_film(self.ctx, C.c_float(offx), C.c_float(offy), C.c_float(n_paths_norm)))

    # ---- photons
    def trace_photons(self, first_index, n_paths, max_depth, radius):
        n = C.c_uint64()
        self._ck(self.lib.bre_trace_photons(self.ctx, C.c_uint64(first_index), C.c_uint64(n_paths),
                                            max_depth, C.c_float(radius), C.byref(n)))
        return n.value

    def upload_photons(self, pos, power, dirs, radius=0.0, radii=None, append=False):
        pos, power, dirs = _f32(pos), _f32(power), _f32(dirs)
        n = pos.size // 3
        r = _f32(radii, (n,)) if radii is not None else None
        self._ck(self.lib.bre_upload_photons(self.ctx, C.c_uint64(n), _fp(pos), _fp(power), _fp(dirs),
                                             C.c_float(radius), _fp(r), int(append)))

    def clear_photons(self):
        self._ck(self.lib.bre_clear_photons(self.ctx))

    def photon_count(self):
        n = C.c_uint64()
        self._ck(self.lib.bre_photon_count(self.ctx, C.byref(n)))
        return n.value

    def download_photons(self):
        n = self.photon_count()
        pos = np.zeros((n, 3), np.float32); pw = np.zeros((n, 3), np.float32)
        dr = np.zeros((n, 3), np.float32); rad = np.zeros(n, np.float32)
        self._ck(self.lib.bre_download_photons(self.ctx, C.c_uint64(n), _fp(pos), _fp(pw), _fp(dr), _fp(rad)))
        return pos, pw, dr, rad

    def save_photons(self, path, n_paths=0):
        self._ck(self.lib.bre_save_photons(self.ctx, os.fsencode(path), C.c_uint64(n_paths)))

    def load_photons(self, path, append=False):
        n, npaths = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.bre_load_photons(self.ctx, os.fsencode(path), int(append), C.byref(n), C.byref(npaths)))
        return n.value, npaths.value

    def film_save(self, path, iterations_done, radius):
        self._ck(self.lib.bre_film_save(self.ctx, os.fsencode(path), C.c_uint64(iterations_done), C.c_float(radius)))

    def film_load(self, path):
        it, r = C.c_uint64(), C.c_float()
        self._ck(self.lib.bre_film_load(self.ctx, os.fsencode(path), C.byref(it), C.byref(r)))
        return it.value, r.value

    def set_radius(self, r):
        self._ck(self.lib.bre_set_radius(self.ctx, C.c_float(r)))

    def trace_exchange_photons(self, n_paths_total, max_depth, radius):
        n = C.c_uint64()
        self._ck(self.lib.bre_trace_exchange_photons(self.ctx, C.c_uint64(n_paths_total), max_depth,
                                                     C.c_float(radius), C.byref(n)))
        return n.value

    def allgather_photons(self):
        self._ck(self.lib.bre_allgather_photons(self.ctx))

    # ---- build
    def build(self):
        self._ck(self.lib.bre_build(self.ctx))

    def last_build_kind(self):
        return self.lib.bre_last_build_kind(self.ctx)

    def download_tree(self):
        n = self.photon_count()
        order = np.zeros(n, np.uint32); codes = np.zeros(n, np.uint32)
        m = max(n - 1, 0)
        lo = np.zeros((m, 3), np.float32); hi = np.zeros((m, 3), np.float32)
        ch = np.zeros((m, 2), np.int32)
        self._ck(self.lib.bre_download_tree(
            self.ctx, order.ctypes.data_as(C.POINTER(C.c_uint32)),
            codes.ctypes.data_as(C.POINTER(C.c_uint32)), _fp(lo), _fp(hi),
            ch.ctypes.data_as(C.POINTER(C.c_int32))))
        return order, codes, lo, hi, ch

    # ---- query
    def query(self, o_tmax, dirs, n_paths_norm, want_hits=False, hit_capacity=None):
        o = _f32(o_tmax); d = _f32(dirs)
        n = o.size // 4
        out = np.zeros((n, 3), np.float32)
        if not want_hits:
            self._ck(self.lib.bre_query(self.ctx, n, _fp(o), _fp(d), C.c_float(n_paths_norm), _fp(out), None))
            return out
        offs = np.zeros(n + 1, np.uint64)
        hd = HitDump(offs.ctypes.data_as(C.POINTER(C.c_uint64)), None, 0, 0)
        self._ck(self.lib.bre_query(self.ctx, n, _fp(o), _fp(d), C.c_float(n_paths_norm), _fp(out), C.byref(hd)))
        total = hd.total
        idx = np.zeros(max(total, 1), np.uint32)
        hd = HitDump(offs.ctypes.data_as(C.POINTER(C.c_uint64)),
                     idx.ctypes.data_as(C.POINTER(C.c_uint32)), total, 0)
        self._ck(self.lib.bre_query(self.ctx, n, _fp(o), _fp(d), C.c_float(n_paths_norm), None, C.byref(hd)))
        return out, offs, idx[:total]

    def set_ray_grid(self, width, height):
        self._ck(self.lib.bre_set_ray_grid(self.ctx, C.c_uint32(int(width)), C.c_uint32(int(height))))

    def set_tile_sharding(self, on=True):
        self._ck(self.lib.bre_set_tile_sharding(self.ctx, int(on)))

    def set_photon_sharding(self, on=True):
        self._ck(self.lib.bre_set_photon_sharding(self.ctx, int(on)))

    def query_d(self, n, rays_o_ptr, rays_d_ptr, n_paths_norm, out_ptr):
        self._ck(self.lib.bre_query_d(self.ctx, n, C.c_void_p(rays_o_ptr), C.c_void_p(rays_d_ptr),
                                      C.c_float(n_paths_norm), C.c_void_p(out_ptr)))

    def query_to_film(self, o_tmax, dirs, n_paths_norm, pixel_idx=None):
        """o_tmax / dirs may be numpy arrays or raw host pointers (ints) with n given by a tuple."""
        o = _f32(o_tmax); d = _f32(dirs)
        p = None
        if pixel_idx is not None:
            pixel_idx = np.ascontiguousarray(pixel_idx, dtype=np.uint32)
            p = pixel_idx.ctypes.data_as(C.POINTER(C.c_uint32))
        self._ck(self.lib.bre_query_to_film(self.ctx, o.size // 4, _fp(o), _fp(d), p, C.c_float(n_paths_norm)))

    def query_to_film_ptr(self, n, o_ptr, d_ptr, n_paths_norm):
        self._ck(self.lib.bre_query_to_film(self.ctx, n, C.cast(o_ptr, C.POINTER(C.c_float)),
                                            C.cast(d_ptr, C.POINTER(C.c_float)), None, C.c_float(n_paths_norm)))

    def stage_rays_ptr(self, n, o_ptr, d_ptr):
        self._ck(self.lib.bre_stage_rays(self.ctx, n, C.cast(o_ptr, C.POINTER(C.c_float)),
                                         C.cast(d_ptr, C.POINTER(C.c_float))))

    def stage_rays(self, o_tmax, dirs):
        self._staged = (_f32(o_tmax), _f32(dirs))       # keep the host arrays alive until the query
        o, d = self._staged
        self._ck(self.lib.bre_stage_rays(self.ctx, o.size // 4, _fp(o), _fp(d)))

    def query_staged_to_film(self, n_paths_norm, pixel_idx=None):
        p = None
        if pixel_idx is not None:
            pixel_idx = np.ascontiguousarray(pixel_idx, dtype=np.uint32)
            p = pixel_idx.ctypes.data_as(C.POINTER(C.c_uint32))
        self._ck(self.lib.bre_query_staged_to_film(self.ctx, p, C.c_float(n_paths_norm)))

    def film_resolve_ptr(self, it, scale, out_ptr):
        self._ck(self.lib.bre_film_resolve(self.ctx, it, C.c_float(scale), C.cast(out_ptr, C.POINTER(C.c_float))))

    def set_timing(self, on=True):
        self._ck(self.lib.bre_set_timing(self.ctx, int(on)))

    def timings(self):
        ms = (C.c_float * 8)()
        self._ck(self.lib.bre_get_timings(self.ctx, ms))
        return dict(trace=ms[0], build=ms[1], tau=ms[2], query=ms[3], film=ms[4], query_a=ms[5], query_b=ms[6])

    def count(self, o_tmax, dirs):
        o = _f32(o_tmax); d = _f32(dirs)
        self._ck(self.lib.bre_count(self.ctx, o.size // 4, _fp(o), _fp(d)))
        return self.counters()

    def count_d(self, n, rays_o_ptr, rays_d_ptr):
        self._ck(self.lib.bre_count_d(self.ctx, n, C.c_void_p(rays_o_ptr), C.c_void_p(rays_d_ptr)))
        return self.counters()

    def counters(self):
        c = Counters()
        self._ck(self.lib.bre_get_counters(self.ctx, C.byref(c)))
        return {k: getattr(c, k) for k, _ in Counters._fields_}

    # ---- film
    def film_create(self, n_pixels):
        self._ck(self.lib.bre_film_create(self.ctx, n_pixels))
        self._npix = n_pixels

    def film_clear(self):
        self._ck(self.lib.bre_film_clear(self.ctx))

    def film_accumulate(self, rgb, pixel_idx=None):
        rgb = _f32(rgb)
        n = rgb.size // 3
        p = None
        if pixel_idx is not None:
            pixel_idx = np.ascontiguousarray(pixel_idx, dtype=np.uint32)
            p = pixel_idx.ctypes.data_as(C.POINTER(C.c_uint32))
        self._ck(self.lib.bre_film_accumulate(self.ctx, n, p, _fp(rgb)))

    def film_accumulate_d(self, n, rgb_ptr, pixel_idx_ptr=None):
        self._ck(self.lib.bre_film_accumulate_d(self.ctx, n, C.c_void_p(pixel_idx_ptr), C.c_void_p(rgb_ptr)))

    def film_resolve(self, it, scale=1.0):
        out = np.zeros((self._npix, 3), np.float32)
        self._ck(self.lib.bre_film_resolve(self.ctx, it, C.c_float(scale), _fp(out)))
        return out

    def film_resolve_static(self, it, scale=1.0, mean_only=False):
        """bre_film_resolve / bre_film_mean without any allocation (static device scratch, <= 2^21 pixels)."""
        out = np.zeros((self._npix, 3), np.float32)
        self._ck(self.lib.bre_film_resolve_static(self.ctx, it, C.c_float(scale), _fp(out), int(mean_only)))
        return out

    def film_gather(self, it, scale=1.0, mean_only=False, want=True):
        """Collective (tile-sharded camera pass): rank 0 passes want=True and gets the image, the others None."""
        out = np.zeros((self._npix, 3), np.float32) if want else None
        self._ck(self.lib.bre_film_gather(self.ctx, it, C.c_float(scale), _fp(out) if want else None, int(mean_only)))
        return out

    def film_gather_ptr(self, it, scale, out_ptr):
        self._ck(self.lib.bre_film_gather(self.ctx, it, C.c_float(scale), C.cast(out_ptr, C.POINTER(C.c_float)), 0))

    def film_mean(self, it):
        out = np.zeros((self._npix, 3), np.float32)
        self._ck(self.lib.bre_film_mean(self.ctx, it, _fp(out)))
        return out


def shard_range(n, n_ranks, rank, lib=None):
    lib = lib or load_library()
    f, c = C.c_uint64(), C.c_uint64()
    lib.bre_shard_range(C.c_uint64(int(n)), int(n_ranks), int(rank), C.byref(f), C.byref(c))
    return f.value, c.value


def tile_owner(tx, ty, tiles_x, n_ranks, lib=None):
    lib = lib or load_library()
    return lib.bre_tile_owner(C.c_uint32(int(tx)), C.c_uint32(int(ty)), C.c_uint32(int(tiles_x)), int(n_ranks))
