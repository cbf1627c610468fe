"""TEST INFRASTRUCTURE: ctypes access to tests/emu_traversal.cpp, a sequential CPU emulation of the
traversal kernel over the product's own QuadNode encoder (aten_b200/csrc/atn_bvh.h).  It gives the CPU
suite a way to check the encoder without a GPU; the product never loads it."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from aten_b200 import scene_io as sio

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_build", "libemu_traversal.so")
SRC = os.path.join(HERE, "emu_traversal.cpp")
DEPS = [SRC, os.path.join(HERE, "..", "aten_b200", "csrc", "atn_bvh.h"), os.path.join(HERE, "..", "aten_b200", "csrc", "atn_math.cuh"),
        os.path.join(HERE, "..", "aten_b200", "csrc", "atn_cache.h")]
_lib = None

STAT_NAMES = ("nodes", "tris", "entries", "entries_filtered", "pops_culled", "max_sp", "node_count", "slots_used",
              "replaced", "not_nested", "stack_bound", "visited_0slot", "visited_1slot", "visited_2slot", "visited_3slot",
              "visited_4slot")


def lib():
    global _lib
    if _lib is None:
        stale = not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in DEPS)
        if stale:
            os.makedirs(os.path.dirname(SO), exist_ok=True)
            subprocess.run(["g++", "-std=c++17", "-O2", "-fopenmp", "-fPIC", "-ffp-contract=off", "-shared", "-o", SO, SRC],
                           check=True)
        _lib = C.CDLL(SO)
    return _lib


def trace_closest(scene: sio.Scene, rays: np.ndarray, t_max=None, leaves_first=False, entry_filter=True):
    """Closest hits of `rays` through encode_wide(scene) + the kernel's per-ray logic; returns (hits, stats dict)."""
    rays = np.ascontiguousarray(rays, dtype=sio.RAY_DTYPE)
    hits = np.zeros(len(rays), dtype=sio.HIT_DTYPE)
    stats = np.zeros(16, np.uint64)
    tm = None if t_max is None else np.ascontiguousarray(t_max, dtype=np.float32)
    blob = scene.blob()
    rc = lib().emu_trace_wide(C.byref(blob), C.c_void_p(rays.ctypes.data), C.c_void_p(tm.ctypes.data) if tm is not None else None,
                              C.c_size_t(len(rays)), C.c_void_p(hits.ctypes.data), C.c_void_p(stats.ctypes.data),
                              C.c_int(1 if leaves_first else 0), C.c_int(1 if entry_filter else 0))
    if rc != 0:
        raise RuntimeError("encode_wide rejected the scene's node lists")
    return hits, {k: int(v) for k, v in zip(STAT_NAMES, stats)}


def top_update_equals_fresh(base: sio.Scene, moved: sio.Scene) -> int:
    """0 when encode_top(moved's list 0 | base's bottom trees) is byte-identical to the tail of encode_wide(moved)."""
    b0, b1 = base.blob(), moved.blob()
    return int(lib().emu_top_update_equals_fresh(C.byref(b0), C.byref(b1)))


KERNEL_BAND = (5e-5, 3.0, 2e-6, 5e-5)      # kBandRel, kBandInv, kBandS, kBandM of atn_kernels.cu


def trace_closest_resolved(scene: sio.Scene, rays: np.ndarray, t_max=None):
    """What the product computes: the front-to-back walk with the kernel's near-tie band, and the reference-order walk
    (the job of k_exact; here the oracle's) for the rays it flags.  Returns (hits, stats) with stats["flagged"]."""
    from tests import oracle_lib as ol
    hits, flags, st = trace_closest_flagged(scene, rays, KERNEL_BAND, t_max)
    idx = np.nonzero(flags)[0]
    if len(idx):
        sub = ol.trace_closest(scene, np.ascontiguousarray(rays[idx]), None if t_max is None else np.ascontiguousarray(np.asarray(t_max)[idx])).copy()
        miss = sub["objid"] < 0
        sub["primid"][miss] = -1; sub["mtrlid"][miss] = -1; sub["meshid"][miss] = -1
        hits[idx] = sub
    st["flagged"] = int(len(idx))
    return hits, st


def trace_closest_flagged(scene: sio.Scene, rays: np.ndarray, band, t_max=None):
    """The kernel's per-ray logic with the near-tie bookkeeping (band = (rel, inv, S, M), see emu_traversal.cpp);
    returns (hits, flags, stats): flags[i] = 1 where the exact reference-order walk has to decide the ray."""
    rays = np.ascontiguousarray(rays, dtype=sio.RAY_DTYPE)
    hits = np.zeros(len(rays), dtype=sio.HIT_DTYPE)
    flags = np.zeros(len(rays), dtype=np.uint8)
    stats = np.zeros(16, np.uint64)
    tm = None if t_max is None else np.ascontiguousarray(t_max, dtype=np.float32)
    b = np.ascontiguousarray(band, dtype=np.float32)
    blob = scene.blob()
    rc = lib().emu_trace_wide_flag(C.byref(blob), C.c_void_p(rays.ctypes.data), C.c_void_p(tm.ctypes.data) if tm is not None else None,
                                   C.c_size_t(len(rays)), C.c_void_p(hits.ctypes.data), C.c_void_p(stats.ctypes.data),
                                   C.c_void_p(b.ctypes.data), C.c_void_p(flags.ctypes.data))
    if rc != 0:
        raise RuntimeError("encode_wide rejected the scene's node lists")
    return hits, flags, {k: int(v) for k, v in zip(STAT_NAMES, stats)}


def encoded_cache_round_trip(scene: sio.Scene, path: str, other: sio.Scene | None = None) -> int:
    """0 when encode -> file -> read -> encode over the cached bottom trees is byte-identical to the fresh encoding (and the
    file is refused for `other` and for a wrong hash)."""
    b0 = scene.blob()
    b1 = other.blob() if other is not None else None
    return int(lib().emu_encoded_cache_round_trip(C.byref(b0), C.c_char_p(os.fsencode(path)), C.byref(b1) if b1 is not None else None))
