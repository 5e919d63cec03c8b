"""Host-side mirror of the reference's driver interface over the libgravsim C-ABI.

The reference has no operator/plugin API: its drivers hand an array-of-structs particle array
plus N / steps / dt to the force + update functions and read the array back
(cuda.cu:145-173, mpi.c:154-257).  ``GravSim`` mirrors exactly that: ``upload`` (records in the
reference's own 28-byte / 56-byte layouts), ``step``, ``accel``, ``energy``, ``download``.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(_HERE)

INTEGRATOR_EULER = 0
INTEGRATOR_LEAPFROG = 1
G_SI = 6.67430e-11          # cuda.cu:11, mpi.c:9
FLAG_TWO_PASS = 2
FLAG_NO_UNIFORM_MASS = 4
FLAG_OVERLAP = 8
FLAG_NCCL_EXCHANGE = 16
FLAG_COMPENSATED = 32
FLAG_DETERMINISTIC = 128
FLAG_FP64_FORCES = 256
NCCL_ID_BYTES = 128

# every symbol include/gravsim.h declares (checked by the CPU tests)
EXPORTED_SYMBOLS = [
    "gs_config_init", "gs_create", "gs_destroy", "gs_nccl_unique_id",
    "gs_upload_aos_f32", "gs_upload_aos_f64", "gs_download_aos_f32", "gs_download_aos_f64",
    "gs_upload_slice_f32", "gs_upload_slice_f64", "gs_download_slice_f32", "gs_download_slice_f64",
    "gs_step", "gs_step_trajectory", "gs_accel", "gs_accel_f64", "gs_energy", "gs_monitor_start", "gs_monitor_read", "gs_potential_energy_noalloc", "gs_sync", "gs_set_dt",
    "gs_get_info", "gs_kernel_launches", "gs_last_cuda_error", "gs_strerror", "gs_version",
    "gs_host_alloc", "gs_host_free",
    "gs_plan_slice", "gs_plan_cta_range", "gs_plan_tile_owners", "gs_plan_tiles",
    "gs_plan_sym_row", "gs_plan_sym_unit", "gs_plan_sym_cta",
    "gs_ic_reference_uniform_f32", "gs_ic_reference_uniform_f64", "gs_ic_plummer_f32",
    "gs_ic_two_galaxy_f32",
    "gs_state_write", "gs_state_read_header", "gs_state_read",
    "gs_dump_positions_cuda_format", "gs_dump_positions_mpi_format",
]


class GravSimError(RuntimeError):
    def __init__(self, code, what, cuda_error=0):
        self.code = code
        self.cuda_error = cuda_error
        msg = lib().gs_strerror(code).decode() if _LIB is not None else str(code)
        extra = f" (cudaError {cuda_error})" if cuda_error else ""
        super().__init__(f"{what}: {msg}{extra}")


class GsConfig(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("n", C.c_int32), ("dt", C.c_double), ("G", C.c_double),
        ("eps", C.c_double), ("integrator", C.c_int32), ("device", C.c_int32),
        ("rank", C.c_int32), ("world_size", C.c_int32), ("nccl_id", C.c_void_p),
        ("stream", C.c_void_p), ("flags", C.c_uint32), ("variant", C.c_int32),
    ]


class GsInfo(C.Structure):
    _fields_ = [
        ("n", C.c_int32), ("n_pad", C.c_int32), ("i_begin", C.c_int32), ("i_end", C.c_int32),
        ("sm_count", C.c_int32), ("grid", C.c_int32), ("i_tile", C.c_int32), ("j_tile", C.c_int32),
        ("stages", C.c_int32), ("variant", C.c_int32), ("regs_per_thread", C.c_int32),
        ("threads", C.c_int32), ("scratch_bytes", C.c_int64), ("peer_exchange", C.c_int32),
        ("uniform_mass", C.c_int32), ("symmetric", C.c_int32), ("reserved", C.c_int32),
    ]


class GsStateHeader(C.Structure):
    _fields_ = [
        ("magic", C.c_char * 8), ("n", C.c_int32), ("record_bytes", C.c_int32), ("step", C.c_int64),
        ("time", C.c_double), ("dt", C.c_double), ("G", C.c_double), ("eps", C.c_double),
        ("reserved", C.c_uint8 * 8),
    ]


def lib_path():
    return os.path.join(_HERE, "lib", "libgravsim.so")


def build(verbose=False):
    """Compile libgravsim.so and the CLI driver in-tree (nvcc, sm_100a)."""
    r = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libgravsim.so failed")
    return lib_path()


_LIB = None


def lib():
    """The loaded CDLL.  Raises if the CUDA library has not been built: there is no fallback."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(libgravsim has no CPU fallback)")
    L = C.CDLL(path)
    L.gs_strerror.restype = C.c_char_p
    L.gs_strerror.argtypes = [C.c_int]
    L.gs_version.restype = C.c_char_p
    L.gs_kernel_launches.restype = C.c_uint64
    L.gs_kernel_launches.argtypes = [C.c_void_p]
    L.gs_last_cuda_error.argtypes = [C.c_void_p]
    L.gs_config_init.argtypes = [C.POINTER(GsConfig)]
    L.gs_config_init.restype = None
    L.gs_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(GsConfig)]
    L.gs_destroy.argtypes = [C.c_void_p]
    L.gs_destroy.restype = None
    for f in ("gs_upload_aos_f32", "gs_upload_aos_f64", "gs_download_aos_f32", "gs_download_aos_f64"):
        getattr(L, f).argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    for f in ("gs_upload_slice_f32", "gs_upload_slice_f64", "gs_download_slice_f32", "gs_download_slice_f64"):
        getattr(L, f).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    L.gs_step.argtypes = [C.c_void_p, C.c_int]
    L.gs_step_trajectory.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    L.gs_accel.argtypes = [C.c_void_p, C.c_void_p]
    L.gs_accel_f64.argtypes = [C.c_void_p, C.c_void_p]
    L.gs_energy.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.gs_monitor_start.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.gs_monitor_read.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    L.gs_sync.argtypes = [C.c_void_p]
    L.gs_potential_energy_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
    L.gs_set_dt.argtypes = [C.c_void_p, C.c_double]
    L.gs_get_info.argtypes = [C.c_void_p, C.POINTER(GsInfo)]
    L.gs_nccl_unique_id.argtypes = [C.c_void_p]
    L.gs_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    L.gs_host_free.argtypes = [C.c_void_p]
    L.gs_plan_slice.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int] + [C.POINTER(C.c_int)] * 3
    L.gs_plan_cta_range.argtypes = [C.c_int64, C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.gs_plan_tile_owners.argtypes = [C.c_int64, C.c_int64, C.c_int, C.c_int64] + [C.POINTER(C.c_int)] * 3
    L.gs_plan_tiles.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.gs_plan_sym_row.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int)] * 3
    L.gs_plan_sym_unit.argtypes = [C.c_int] * 6 + [C.POINTER(C.c_int)]
    L.gs_plan_sym_cta.argtypes = [C.c_int] * 6 + [C.POINTER(C.c_int)] * 3
    L.gs_ic_reference_uniform_f32.argtypes = [C.c_void_p, C.c_int, C.c_uint64]
    L.gs_ic_reference_uniform_f64.argtypes = [C.c_void_p, C.c_int, C.c_uint64]
    L.gs_ic_plummer_f32.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_double]
    L.gs_ic_two_galaxy_f32.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_double, C.c_double, C.c_double]
    L.gs_state_write.argtypes = [C.c_char_p, C.POINTER(GsStateHeader), C.c_void_p]
    L.gs_state_read_header.argtypes = [C.c_char_p, C.POINTER(GsStateHeader)]
    L.gs_state_read.argtypes = [C.c_char_p, C.POINTER(GsStateHeader), C.c_void_p, C.c_size_t]
    L.gs_dump_positions_cuda_format.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    L.gs_dump_positions_mpi_format.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    _LIB = L
    return L


def _check(code, what, ctx=None):
    if code != 0:
        ce = lib().gs_last_cuda_error(ctx) if ctx else 0
        raise GravSimError(code, what, ce)


def _records(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    if a.ndim != 2 or a.shape[1] != 7:
        raise ValueError("particle records are (n, 7): pos[3], vel[3], mass")
    return a


# ---- host utilities -----------------------------------------------------------------------------
def ic_reference_uniform(n, seed=1234, dtype=np.float32):
    """Sun/Earth/Mars + uniform random bodies (cuda.cu:82-93,129-131; mpi.c:76-104), seeded."""
    out = np.empty((n, 7), dtype=dtype)
    fn = lib().gs_ic_reference_uniform_f32 if dtype == np.float32 else lib().gs_ic_reference_uniform_f64
    _check(fn(out.ctypes.data, n, seed), "gs_ic_reference_uniform")
    return out


def ic_plummer(n, seed=1234, rmax=10.0):
    out = np.empty((n, 7), dtype=np.float32)
    _check(lib().gs_ic_plummer_f32(out.ctypes.data, n, seed, rmax), "gs_ic_plummer_f32")
    return out


def ic_two_galaxy(n, seed=1234, sep=8.0, impact=2.0, vrel=0.5):
    out = np.empty((n, 7), dtype=np.float32)
    _check(lib().gs_ic_two_galaxy_f32(out.ctypes.data, n, seed, sep, impact, vrel), "gs_ic_two_galaxy_f32")
    return out


def plan_tiles(variant=0):
    it, jt = C.c_int(), C.c_int()
    _check(lib().gs_plan_tiles(variant, C.byref(it), C.byref(jt)), "gs_plan_tiles")
    return it.value, jt.value


def plan_slice(n, world_size, rank, variant=0):
    a, b, c = C.c_int(), C.c_int(), C.c_int()
    _check(lib().gs_plan_slice(n, world_size, rank, variant, C.byref(a), C.byref(b), C.byref(c)), "gs_plan_slice")
    return a.value, b.value, c.value


def plan_cta_range(ni_tiles, nj_tiles, grid, cta):
    a, b = C.c_int64(), C.c_int64()
    _check(lib().gs_plan_cta_range(ni_tiles, nj_tiles, grid, cta, C.byref(a), C.byref(b)), "gs_plan_cta_range")
    return a.value, b.value


def plan_tile_owners(ni_tiles, nj_tiles, grid, it):
    a, b, s = C.c_int(), C.c_int(), C.c_int()
    _check(lib().gs_plan_tile_owners(ni_tiles, nj_tiles, grid, it, C.byref(a), C.byref(b), C.byref(s)),
           "gs_plan_tile_owners")
    return a.value, b.value, bool(s.value)


def plan_sym_row(n, world_size, rank, variant, row):
    a, b, c = C.c_int(), C.c_int(), C.c_int()
    _check(lib().gs_plan_sym_row(n, world_size, rank, variant, row, C.byref(a), C.byref(b), C.byref(c)), "gs_plan_sym_row")
    return a.value, b.value, c.value          # rows of the rank, units of this row, leading plain units


def plan_sym_unit(n, world_size, rank, variant, row, k):
    j = C.c_int()
    _check(lib().gs_plan_sym_unit(n, world_size, rank, variant, row, k, C.byref(j)), "gs_plan_sym_unit")
    return j.value


def plan_sym_cta(n, world_size, rank, variant, grid, cta):
    a, b, c = C.c_int(), C.c_int(), C.c_int()
    _check(lib().gs_plan_sym_cta(n, world_size, rank, variant, grid, cta, C.byref(a), C.byref(b), C.byref(c)), "gs_plan_sym_cta")
    return a.value, b.value, c.value


def nccl_unique_id():
    buf = (C.c_ubyte * NCCL_ID_BYTES)()
    _check(lib().gs_nccl_unique_id(buf), "gs_nccl_unique_id")
    return bytes(buf)


def state_write(path, records, step=0, time=0.0, dt=0.0, G=G_SI, eps=0.0):
    records = np.ascontiguousarray(records)
    h = GsStateHeader()
    h.n = records.shape[0]
    h.record_bytes = 28 if records.dtype == np.float32 else 56
    h.step, h.time, h.dt, h.G, h.eps = step, time, dt, G, eps
    _check(lib().gs_state_write(path.encode(), C.byref(h), records.ctypes.data), "gs_state_write")


def state_read(path):
    h = GsStateHeader()
    _check(lib().gs_state_read_header(path.encode(), C.byref(h)), "gs_state_read_header")
    out = np.empty((h.n, 7), dtype=np.float32 if h.record_bytes == 28 else np.float64)
    _check(lib().gs_state_read(path.encode(), C.byref(h), out.ctypes.data, out.nbytes), "gs_state_read")
    return out, h


def _libc_file(path, mode):
    libc = C.CDLL(None)
    libc.fopen.restype = C.c_void_p
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [C.c_void_p]
    f = libc.fopen(path.encode(), mode.encode())
    if not f:
        raise OSError(f"cannot open {path}")
    return libc, f


def dump_positions_cuda_format(path, records28, print_particles=10):
    r = _records(records28, np.float32)
    libc, f = _libc_file(path, "w")
    try:
        _check(lib().gs_dump_positions_cuda_format(f, r.ctypes.data, r.shape[0], print_particles), "dump")
    finally:
        libc.fclose(f)


def dump_positions_mpi_format(path, records56):
    r = _records(records56, np.float64)
    libc, f = _libc_file(path, "w")
    try:
        _check(lib().gs_dump_positions_mpi_format(f, r.ctypes.data, r.shape[0]), "dump")
    finally:
        libc.fclose(f)


NOALLOC_MAX_BODIES = 4194304


def potential_energy_noalloc(pos_packed_ptr, n_pad, G, eps, potential_ptr, stream=0):
    """gs_potential_energy_noalloc on caller-owned device memory (raw pointers); allocates nothing.
    Enqueues on `stream` and returns; read *potential_ptr after synchronising that stream."""
    _check(lib().gs_potential_energy_noalloc(C.c_void_p(pos_packed_ptr), int(n_pad), float(G), float(eps),
                                             C.c_void_p(potential_ptr), C.c_void_p(stream)),
           "gs_potential_energy_noalloc")


# ---- the engine -----------------------------------------------------------------------------------
class GravSim:
    """One GPU's engine context (gs_ctx).  Mirrors the reference drivers' use of the hot path:

        sim = GravSim(n, dt=3600.0)            # cuda.cu:121-123 / mpi.c:146-148 constants
        sim.upload(particles)                  # (n,7) float32 [cuda.cu Particle] or float64 [mpi.c Particle]
        sim.step(500)                          # cuda.cu:155-167 / mpi.c:191-237
        out = sim.download()                   # cuda.cu:173 / mpi.c:250-257
    """

    def __init__(self, n, dt=3600.0, G=G_SI, eps=0.0, integrator=INTEGRATOR_EULER, device=-1,
                 rank=0, world_size=1, nccl_id=None, stream=None, flags=0, variant=0):
        L = lib()
        cfg = GsConfig()
        L.gs_config_init(C.byref(cfg))
        cfg.n, cfg.dt, cfg.G, cfg.eps = int(n), float(dt), float(G), float(eps)
        cfg.integrator, cfg.device, cfg.rank, cfg.world_size = integrator, device, rank, world_size
        cfg.flags, cfg.variant = flags, variant
        self._id_buf = None
        if nccl_id is not None:
            self._id_buf = (C.c_ubyte * NCCL_ID_BYTES).from_buffer_copy(nccl_id)
            cfg.nccl_id = C.cast(self._id_buf, C.c_void_p)
        if stream is not None:
            cfg.stream = C.c_void_p(int(stream))
        self._ctx = C.c_void_p()
        self.n = int(n)
        _check(L.gs_create(C.byref(self._ctx), C.byref(cfg)), "gs_create")
        self.info = GsInfo()
        _check(L.gs_get_info(self._ctx, C.byref(self.info)), "gs_get_info", self._ctx)

    # -- data in/out
    def upload(self, particles):
        particles = np.asarray(particles)
        if particles.dtype == np.float64:
            r = _records(particles, np.float64)
            _check(lib().gs_upload_aos_f64(self._ctx, r.ctypes.data, r.shape[0]), "gs_upload_aos_f64", self._ctx)
        else:
            r = _records(particles, np.float32)
            _check(lib().gs_upload_aos_f32(self._ctx, r.ctypes.data, r.shape[0]), "gs_upload_aos_f32", self._ctx)
        lib().gs_get_info(self._ctx, C.byref(self.info))

    def upload_ptr(self, ptr, n, record_bytes=28):
        fn = lib().gs_upload_aos_f32 if record_bytes == 28 else lib().gs_upload_aos_f64
        _check(fn(self._ctx, C.c_void_p(ptr), n), "gs_upload_aos", self._ctx)
        lib().gs_get_info(self._ctx, C.byref(self.info))

    def download(self, dtype=np.float32):
        out = np.empty((self.n, 7), dtype=dtype)
        fn = lib().gs_download_aos_f32 if dtype == np.float32 else lib().gs_download_aos_f64
        _check(fn(self._ctx, out.ctypes.data, self.n), "gs_download_aos", self._ctx)
        return out

    def download_ptr(self, ptr, n, record_bytes=28):
        fn = lib().gs_download_aos_f32 if record_bytes == 28 else lib().gs_download_aos_f64
        _check(fn(self._ctx, C.c_void_p(ptr), n), "gs_download_aos", self._ctx)

    # -- this rank's own records only (mpi.c:184-187's slice)
    @property
    def my_range(self):
        return self.info.i_begin, self.info.i_end

    def upload_slice(self, my_particles):
        """Collective: every rank passes the records of ITS bodies [i_begin, i_end) only."""
        a = np.asarray(my_particles)
        dt = np.float64 if a.dtype == np.float64 else np.float32
        r = np.ascontiguousarray(a, dtype=dt).reshape(-1, 7)
        self.upload_slice_ptr(r.ctypes.data, r.shape[0], 56 if dt == np.float64 else 28)

    def upload_slice_ptr(self, ptr, count, record_bytes=28):
        fn = lib().gs_upload_slice_f32 if record_bytes == 28 else lib().gs_upload_slice_f64
        _check(fn(self._ctx, C.c_void_p(ptr), self.info.i_begin, count), "gs_upload_slice", self._ctx)
        lib().gs_get_info(self._ctx, C.byref(self.info))

    def download_slice(self, dtype=np.float32):
        lo, hi = self.my_range
        out = np.empty((max(hi - lo, 0), 7), dtype=dtype)
        self.download_slice_ptr(out.ctypes.data, out.shape[0], 28 if dtype == np.float32 else 56)
        return out

    def download_slice_ptr(self, ptr, count, record_bytes=28):
        fn = lib().gs_download_slice_f32 if record_bytes == 28 else lib().gs_download_slice_f64
        _check(fn(self._ctx, C.c_void_p(ptr), self.info.i_begin, count), "gs_download_slice", self._ctx)

    # -- hot path
    def step(self, nsteps=1):
        _check(lib().gs_step(self._ctx, int(nsteps)), "gs_step", self._ctx)

    def step_trajectory(self, nsteps, every=1):
        """Like step(), returning positions recorded every `every` steps: (nsteps // every, n, 3) float32
        (what pyspark.py:104-121 collects as `trajectories`)."""
        count = int(nsteps) // int(every)
        out = np.empty((count, self.n, 3), dtype=np.float32)
        got = C.c_int()
        _check(lib().gs_step_trajectory(self._ctx, int(nsteps), int(every), out.ctypes.data, count, C.byref(got)),
               "gs_step_trajectory", self._ctx)
        return out[:got.value]

    def accel(self):
        """Accelerations (n_local, 3) float32 of this rank's bodies on the current snapshot."""
        n_local = self.info.i_end - self.info.i_begin
        out = np.empty((max(n_local, 0), 3), dtype=np.float32)
        _check(lib().gs_accel(self._ctx, out.ctypes.data), "gs_accel", self._ctx)
        return out

    def accel_f64(self):
        """The same as float64 (hi + lo words: beyond fp32 only with FLAG_FP64_FORCES)."""
        n_local = self.info.i_end - self.info.i_begin
        out = np.empty((max(n_local, 0), 3), dtype=np.float64)
        _check(lib().gs_accel_f64(self._ctx, out.ctypes.data), "gs_accel_f64", self._ctx)
        return out

    def energy(self):
        ke, pe = C.c_double(), C.c_double()
        mom = (C.c_double * 3)()
        _check(lib().gs_energy(self._ctx, C.byref(ke), C.byref(pe), mom), "gs_energy", self._ctx)
        return ke.value, pe.value, np.array(list(mom))

    def monitor_start(self, every, capacity=1024):
        """Sample energy and momentum on the device after every `every`-th step (0 = off)."""
        _check(lib().gs_monitor_start(self._ctx, int(every), int(capacity)), "gs_monitor_start", self._ctx)
        self._mon_capacity = int(capacity)

    def monitor_read(self):
        """(k, 8) float64: step, time, KE, PE, px, py, pz, 0 of the samples taken so far."""
        cap = getattr(self, "_mon_capacity", 0)
        out = np.zeros((max(cap, 1), 8))
        got = C.c_int()
        _check(lib().gs_monitor_read(self._ctx, out.ctypes.data, cap, C.byref(got)), "gs_monitor_read", self._ctx)
        return out[:got.value]

    def sync(self):
        _check(lib().gs_sync(self._ctx), "gs_sync", self._ctx)

    def set_dt(self, dt):
        _check(lib().gs_set_dt(self._ctx, float(dt)), "gs_set_dt", self._ctx)

    @property
    def kernel_launches(self):
        return int(lib().gs_kernel_launches(self._ctx))

    def close(self):
        if self._ctx:
            lib().gs_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
