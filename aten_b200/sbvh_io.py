"""Reader / writer of the reference's `.sbvh` tree files.

On-disk layout written by ``sbvh::exportTree`` and read by ``sbvh::importTree``
(reference src/libaten/accelerator/sbvh.cpp:1220-1436), little endian::

    SbvhFileHeader (44 B): char magic[4] = "SBVH"; u8 version[4]; u32 nodeNum; u32 maxDepth;
                           u32 cntMtrlForVoxel; float boxmin[3]; float boxmax[3]
    cntMtrlForVoxel x { i32 mtrlid; i32 alignedLen; char name[alignedLen] }   (voxel-LOD material names)
    nodeNum x ThreadedSbvhNode (48 B)  == AtnBvhNode: boxmin, hit, boxmax, miss, isleaf, triid, voxeldepth, mtrlid

``read_sbvh`` applies what ``importTree`` applies: ``triid += offset_tri_idx`` for leaves and the
re-mapping of voxel material ids through the name table.  The node array can be used directly as one
bottom-layer list of an ``AtnSceneBlob`` (scene_io.Scene.node_lists[k]) -- the re-encoding into the
traversal layout happens in ``atn_upload_scene``.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import Callable, Dict, Optional

import numpy as np

from .scene_io import NODE_DTYPE

HEADER = struct.Struct("<4s4BIII3f3f")
assert HEADER.size == 44


@dataclass
class SbvhFile:
    version: tuple
    max_depth: int
    boxmin: np.ndarray
    boxmax: np.ndarray
    voxel_materials: Dict[int, str]
    nodes: np.ndarray          # NODE_DTYPE


def read_sbvh(path: str, offset_tri_idx: int = 0,
              material_index_by_name: Optional[Callable[[str], int]] = None) -> SbvhFile:
    with open(path, "rb") as fp:
        buf = fp.read()
    if len(buf) < HEADER.size:
        raise ValueError(f"{path}: too short for an SBVH header")
    f = HEADER.unpack_from(buf, 0)
    magic, version, node_num, max_depth, n_mtrl = f[0], f[1:5], f[5], f[6], f[7]
    if magic != b"SBVH":
        raise ValueError(f"{path}: bad magic {magic!r}")
    boxmin = np.array(f[8:11], np.float32)
    boxmax = np.array(f[11:14], np.float32)
    off = HEADER.size
    names: Dict[int, str] = {}
    for _ in range(n_mtrl):
        mtrlid, alen = struct.unpack_from("<ii", buf, off)
        off += 8
        if not (0 < alen < 128):
            raise ValueError(f"{path}: bad material name length {alen}")
        names[mtrlid] = buf[off:off + alen].split(b"\0", 1)[0].decode("utf-8", "replace")
        off += alen
    need = node_num * NODE_DTYPE.itemsize
    if len(buf) - off < need:
        raise ValueError(f"{path}: truncated ({len(buf) - off} bytes of nodes, {need} expected)")
    nodes = np.frombuffer(buf, NODE_DTYPE, node_num, off).copy()
    # importTree (sbvh.cpp:1401-1424)
    if offset_tri_idx > 0 or names:
        leaf = nodes["primid"] >= 0
        nodes["primid"][leaf] += np.float32(offset_tri_idx)
        if names:
            has = nodes["meshid"] >= 0            # ThreadedSbvhNode::mtrlid sits where meshid is
            for i in np.nonzero(has)[0]:
                old = int(nodes["meshid"][i])
                if old not in names:
                    raise ValueError(f"{path}: node {i} refers to voxel material {old} that is not in the name table")
                if material_index_by_name is not None:
                    new = material_index_by_name(names[old])
                    if new < 0:
                        raise ValueError(f"{path}: material {names[old]!r} not found in the scene")
                    nodes["meshid"][i] = np.float32(new)
    return SbvhFile(tuple(version), int(max_depth), boxmin, boxmax, names, nodes)


def write_sbvh(path: str, nodes: np.ndarray, max_depth: int, boxmin, boxmax,
               voxel_materials: Optional[Dict[int, str]] = None, version=(0, 0, 0, 1)) -> None:
    """``sbvh::exportTree`` layout (sbvh.cpp:1236-1340)."""
    nodes = np.ascontiguousarray(nodes, NODE_DTYPE)
    voxel_materials = voxel_materials or {}
    with open(path, "wb") as fp:
        fp.write(HEADER.pack(b"SBVH", *version, len(nodes), max_depth, len(voxel_materials),
                             *[float(x) for x in boxmin], *[float(x) for x in boxmax]))
        for mtrlid in sorted(voxel_materials):
            raw = voxel_materials[mtrlid].encode("utf-8")
            alen = (len(raw) + 3) // 4 * 4
            fp.write(struct.pack("<ii", mtrlid, alen))
            fp.write(raw + b"\0" * (alen - len(raw)))
        fp.write(nodes.tobytes())
