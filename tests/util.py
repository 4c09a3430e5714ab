"""ctypes wrappers of the CPU checkers used by the tests.

Oracle  -> oracle/liboracle.so           (plain C++ restatement)
RefOracle -> oracle/_ref/liboracle_ref.so (the reference's own code, host build)
HostSim -> tests/hostsim/libhostsim.so   (the kernels' per-thread logic on the host)
All three are test infrastructure; the product never loads them.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
f32 = np.float32


def fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def _build_oracle():
    so = ROOT / "oracle" / "liboracle.so"
    srcs = list((ROOT / "oracle").glob("oracle_*.cpp")) + list((ROOT / "oracle").glob("*.h"))
    if not so.exists() or any(s.stat().st_mtime > so.stat().st_mtime for s in srcs):
        r = subprocess.run(["make", "-C", str(ROOT / "oracle"), "liboracle.so"], capture_output=True, text=True)
        if r.returncode != 0 and not so.exists():
            raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    return so


def _build_hostsim():
    d = ROOT / "tests" / "hostsim"
    so = d / "libhostsim.so"
    srcs = list(d.glob("*.cpp")) + list((ROOT / "optixpathtracer_b200" / "csrc").glob("*.cuh"))
    if not so.exists() or any(s.stat().st_mtime > so.stat().st_mtime for s in srcs):
        cmd = ["g++", "-std=c++17", "-O2", "-fPIC", "-fopenmp", "-ffp-contract=off", "-shared",
               "-I/usr/local/cuda/include", "-o", str(so)] + [str(p) for p in sorted(d.glob("*.cpp"))]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0 and not so.exists():
            raise RuntimeError(r.stderr)
    return so


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(str(_build_oracle()))

    def closest(self, soup, org, dirs, tmin=0.0, tmax=100.0):
        soup = np.ascontiguousarray(soup, dtype=f32).reshape(-1, 9)
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = len(o)
        prim = np.empty(n, np.int32)
        t, u, v = (np.empty(n, f32) for _ in range(3))
        bf = np.empty(n, np.uint8)
        self.lib.oracle_closest_bruteforce(fp(soup), C.c_int64(len(soup)), fp(o), fp(d), C.c_int64(n),
                                           C.c_float(tmin), C.c_float(tmax), vp(prim), fp(t), fp(u),
                                           fp(v), vp(bf))
        return prim, t, u, v, bf

    def any(self, soup, org, dirs, tmax, tmin=0.0):
        soup = np.ascontiguousarray(soup, dtype=f32).reshape(-1, 9)
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = len(o)
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=f32), (n,)), dtype=f32)
        occ = np.empty(n, np.uint8)
        self.lib.oracle_any_bruteforce(fp(soup), C.c_int64(len(soup)), fp(o), fp(d), C.c_int64(n),
                                       C.c_float(tmin), fp(tm), vp(occ))
        return occ

    def tie_count(self, soup, org, dirs, prim, tmin=0.0, tmax=100.0):
        soup = np.ascontiguousarray(soup, dtype=f32).reshape(-1, 9)
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        prim = np.ascontiguousarray(prim, dtype=np.int32)
        ties = np.empty(len(o), np.int32)
        self.lib.oracle_tie_count(fp(soup), C.c_int64(len(soup)), fp(o), fp(d), C.c_int64(len(o)),
                                  C.c_float(tmin), C.c_float(tmax), vp(prim), vp(ties))
        return ties


class HostSim:
    def __init__(self):
        self.lib = C.CDLL(str(_build_hostsim()))

    def build(self, soup):
        soup = np.ascontiguousarray(soup, dtype=f32).reshape(-1, 9)
        self._soup = soup
        n = self.lib.hostsim_build(fp(soup), len(soup))
        return n, self.lib.hostsim_check_topology()

    def closest(self, org, dirs, tmin=0.0, tmax=100.0):
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = len(o)
        prim = np.empty(n, np.int32)
        t, u, v = (np.empty(n, f32) for _ in range(3))
        bf = np.empty(n, np.uint8)
        cnt = (C.c_uint64 * 3)()
        self.lib.hostsim_trace_closest(fp(o), fp(d), C.c_int64(n), C.c_float(tmin), C.c_float(tmax),
                                       vp(prim), fp(t), fp(u), fp(v), vp(bf), cnt)
        return prim, t, u, v, bf, tuple(cnt)

    def any(self, org, dirs, tmax, tmin=0.0):
        o = np.ascontiguousarray(org, dtype=f32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=f32).reshape(-1, 3)
        n = len(o)
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=f32), (n,)), dtype=f32)
        occ = np.empty(n, np.uint8)
        self.lib.hostsim_trace_any(fp(o), fp(d), C.c_int64(n), C.c_float(tmin), fp(tm), vp(occ))
        return occ


def bits(a):
    return np.ascontiguousarray(a, dtype=f32).view(np.uint32)


def random_rays(rng, n, lo, hi):
    o = rng.uniform(lo, hi, (n, 3)).astype(f32)
    d = rng.normal(size=(n, 3))
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    return o, d
