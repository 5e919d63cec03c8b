"""Shared fixtures: the package under its import name, the oracle (checker) libraries, markers."""
import ctypes as C
import importlib.util
import os
import subprocess
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(REPO, "gravity-simulator-using-mpi-spark-and-cuda_b200")
ORACLE_DIR = os.path.join(REPO, "oracle")
GOLDEN_DIR = os.path.join(REPO, "tests", "golden")
G_SI = 6.67430e-11


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def load_package():
    if "gravsim_b200" in sys.modules:
        return sys.modules["gravsim_b200"]
    spec = importlib.util.spec_from_file_location(
        "gravsim_b200", os.path.join(PKG_DIR, "__init__.py"), submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["gravsim_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def gs():
    mod = load_package()
    if not os.path.exists(mod.lib_path()):
        mod.build()
    return mod


_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)


class Oracle:
    """ctypes view of oracle/libgsoracle.so (our fp64 restatement).  TEST-ONLY."""

    def __init__(self):
        path = os.path.join(ORACLE_DIR, "libgsoracle.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-C", ORACLE_DIR, "port"], check=True, capture_output=True)
        self.lib = C.CDLL(path)
        L = self.lib
        L.gso_accel.argtypes = [_dp, C.c_int, C.c_double, C.c_double, _dp]
        L.gso_accel_subset.argtypes = [_dp, C.c_int, C.c_double, C.c_double, _ip, C.c_int, _dp]
        L.gso_accel_massfree_subset.argtypes = [_dp, C.c_int, C.c_double, C.c_double, _ip, C.c_int, _dp, _dp]
        L.gso_step_sync.argtypes = [_dp, C.c_int, C.c_double, C.c_double, C.c_double]
        L.gso_step_kdk.argtypes = [_dp, C.c_int, C.c_double, C.c_double, C.c_double]
        L.gso_step_aswritten.argtypes = [_dp, C.c_int, C.c_double, C.c_double, C.c_int]
        L.gso_run_sync.argtypes = [_dp, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
        L.gso_run_kdk.argtypes = [_dp, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int]
        L.gso_run_aswritten.argtypes = [_dp, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int]
        L.gso_energy.argtypes = [_dp, C.c_int, C.c_double, C.c_double, _dp, _dp, _dp]
        L.gso_cuda_literal_forces_f32.argtypes = [_fp, C.c_int, _fp]
        L.gso_cuda_literal_update_f32.argtypes = [_fp, C.c_int, _fp, C.c_float]

    @staticmethod
    def _d(a):
        return a.ctypes.data_as(_dp)

    def accel(self, p, G=G_SI, eps=0.0):
        p = np.ascontiguousarray(p, dtype=np.float64)
        out = np.zeros((p.shape[0], 3))
        self.lib.gso_accel(self._d(p), p.shape[0], G, eps, self._d(out))
        return out

    def accel_subset(self, p, idx, G=G_SI, eps=0.0, massfree=False):
        p = np.ascontiguousarray(p, dtype=np.float64)
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        out = np.zeros((idx.size, 3))
        if massfree:
            ns = np.zeros(idx.size)
            self.lib.gso_accel_massfree_subset(self._d(p), p.shape[0], G, eps, idx.ctypes.data_as(_ip), idx.size,
                                               self._d(out), self._d(ns))
            return out, ns
        self.lib.gso_accel_subset(self._d(p), p.shape[0], G, eps, idx.ctypes.data_as(_ip), idx.size, self._d(out))
        return out

    def run_sync(self, p, dt, steps, G=G_SI, eps=0.0):
        q = np.array(p, dtype=np.float64, order="C")
        self.lib.gso_run_sync(self._d(q), q.shape[0], G, eps, dt, steps)
        return q

    def run_kdk(self, p, dt, steps, G=G_SI, eps=0.0):
        q = np.array(p, dtype=np.float64, order="C")
        self.lib.gso_run_kdk(self._d(q), q.shape[0], G, eps, dt, steps)
        return q

    def run_aswritten(self, p, dt, steps, np_ranks, G=G_SI):
        q = np.array(p, dtype=np.float64, order="C")
        self.lib.gso_run_aswritten(self._d(q), q.shape[0], G, dt, steps, np_ranks)
        return q

    def energy(self, p, G=G_SI, eps=0.0):
        p = np.ascontiguousarray(p, dtype=np.float64)
        ke, pe = C.c_double(), C.c_double()
        mom = np.zeros(3)
        self.lib.gso_energy(self._d(p), p.shape[0], G, eps, C.byref(ke), C.byref(pe), self._d(mom))
        return ke.value, pe.value, mom

    def cuda_literal_forces(self, p28):
        p28 = np.ascontiguousarray(p28, dtype=np.float32)
        f = np.zeros((p28.shape[0], 3), dtype=np.float32)
        self.lib.gso_cuda_literal_forces_f32(p28.ctypes.data_as(_fp), p28.shape[0], f.ctypes.data_as(_fp))
        return f


class Reference:
    """The reference's own compiled code (oracle/_ref/, built from /root/reference).  TEST-ONLY."""

    def __init__(self):
        self.dir = os.path.join(ORACLE_DIR, "_ref")
        if not os.path.exists(os.path.join(self.dir, "libgsref.so")) and os.path.exists("/root/reference/mpi.c"):
            subprocess.run(["make", "-C", ORACLE_DIR, "ref"], check=False, capture_output=True)
        so = os.path.join(self.dir, "libgsref.so")
        self.available = os.path.exists(so) and os.path.exists(os.path.join(self.dir, "mpi_ref"))
        if self.available:
            self.lib = C.CDLL(so)
            self.lib.gsref_accel.argtypes = [_dp, C.c_int, _dp]
            self.lib.gsref_accel_subset.argtypes = [_dp, C.c_int, _ip, C.c_int, _dp]
            self.lib.gsref_step_sync.argtypes = [_dp, C.c_int, C.c_double]

    def accel(self, p):
        p = np.ascontiguousarray(p, dtype=np.float64)
        out = np.zeros((p.shape[0], 3))
        self.lib.gsref_accel(p.ctypes.data_as(_dp), p.shape[0], out.ctypes.data_as(_dp))
        return out

    def step_sync(self, p, dt, steps=1):
        q = np.array(p, dtype=np.float64, order="C")
        for _ in range(steps):
            self.lib.gsref_step_sync(q.ctypes.data_as(_dp), q.shape[0], dt)
        return q

    def run_main(self, p, dt, steps, np_ranks, workdir, binary="mpi_ref"):
        """Run the reference's own main() (mpi.c:140-269) on the given bodies; returns (state, seconds)."""
        p = np.ascontiguousarray(p, dtype=np.float64)
        sin, sout = os.path.join(workdir, "state_in.bin"), os.path.join(workdir, "state_out.bin")
        p.tofile(sin)
        env = dict(os.environ, GS_MPI_NP=str(np_ranks), GS_REF_N=str(p.shape[0]), GS_REF_STEPS=str(steps),
                   GS_REF_DT=repr(float(dt)), GS_REF_STATE=sin, GS_REF_OUT=sout)
        r = subprocess.run([os.path.join(self.dir, binary)], env=env, cwd=workdir, capture_output=True, text=True,
                           timeout=600)
        assert r.returncode == 0, r.stderr
        span = [l for l in r.stdout.splitlines() if l.startswith("GS_REF_WTIME_SPAN")]
        return np.fromfile(sout).reshape(-1, 7), float(span[-1].split()[1])


@pytest.fixture(scope="session")
def oracle():
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    return Reference()


def floor_err(a, ref, normsum):
    """per-body |a - ref| / max(|ref|, 1e-3 * sum_j |a_ij|): the relative error with the floor SURVEY.md section 7
    asks for where |a_i| is a near-total cancellation (centre of a Plummer sphere).  Compared with 1e-5."""
    num = np.linalg.norm(np.asarray(a, dtype=np.float64) - ref, axis=1)
    return num / np.maximum(np.linalg.norm(ref, axis=1), 1e-3 * normsum)


def rel_err(a, ref):
    """per-body |a - ref| / |ref| (3-vectors)."""
    num = np.linalg.norm(np.asarray(a, dtype=np.float64) - ref, axis=1)
    den = np.linalg.norm(ref, axis=1)
    return num / np.where(den > 0, den, 1.0)
