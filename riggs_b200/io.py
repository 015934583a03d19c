"""On-disk formats of the reference, so that its checkpoints can be loaded into (and written from) this package
(SURVEY section 8f, N4): the Gaussian point cloud ``point_cloud.ply`` (scene/gaussian_model.py:237-336), the skeleton
tree ``skeleton_tree.npz`` (train_rig.py:233, 245) and the skeleton module's ``skeleton.pth`` state dict
(scene/skeleton_model.py:46, 66).

The reference reads and writes the PLY through the ``plyfile`` package; this is a plain numpy reader / writer of the
same files (binary little-endian or ASCII, float properties, one ``vertex`` element).  Host-side code: nothing here
touches the GPU.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import numpy as np
import torch

_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2",
              "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}


def construct_list_of_attributes(n_dc: int = 3, n_rest: int = 45, n_scale: int = 3, n_rot: int = 4, fea_dim: int = 0) -> List[str]:
    """Property order of the reference's PLY (scene/gaussian_model.py:237-251)."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(n_dc)]
    names += [f"f_rest_{i}" for i in range(n_rest)]
    names.append("opacity")
    names += [f"scale_{i}" for i in range(n_scale)]
    names += [f"rot_{i}" for i in range(n_rot)]
    names += [f"fea_{i}" for i in range(fea_dim)]
    return names


def save_ply(path: str, xyz, features_dc, features_rest, opacity, scaling, rotation, feature=None) -> None:
    """GaussianModel.save_ply (:253-276).  features_dc (P,1,3) and features_rest (P,M-1,3) are stored channel-major
    (``transpose(1, 2).flatten(1)``), normals as zeros; binary little-endian float32 like plyfile's default."""
    to_np = lambda t: t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)
    xyz = to_np(xyz).astype(np.float32)
    P = xyz.shape[0]
    f_dc = to_np(features_dc).astype(np.float32).transpose(0, 2, 1).reshape(P, -1)
    f_rest = to_np(features_rest).astype(np.float32).transpose(0, 2, 1).reshape(P, -1)
    cols = [xyz, np.zeros_like(xyz), f_dc, f_rest, to_np(opacity).astype(np.float32).reshape(P, -1),
            to_np(scaling).astype(np.float32).reshape(P, -1), to_np(rotation).astype(np.float32).reshape(P, -1)]
    fea_dim = 0
    if feature is not None:
        fea = to_np(feature).astype(np.float32).reshape(P, -1)
        fea_dim = fea.shape[1]
        cols.append(fea)
    names = construct_list_of_attributes(f_dc.shape[1], f_rest.shape[1], cols[5].shape[1], cols[6].shape[1], fea_dim)
    table = np.ascontiguousarray(np.concatenate(cols, axis=1), dtype="<f4")
    assert table.shape[1] == len(names)
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    header = ["ply", "format binary_little_endian 1.0", f"element vertex {P}"] + [f"property float {n}" for n in names] + ["end_header"]
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(table.tobytes())


def read_ply_vertices(path: str) -> Dict[str, np.ndarray]:
    """{property name: (P,) array} of the first element of a PLY file (binary little-endian or ASCII)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, count, props, in_first, n_elements = None, 0, [], False, 0
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: truncated PLY header")
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                n_elements += 1
                in_first = n_elements == 1
                if in_first:
                    count = int(tok[2])
            elif tok[0] == "property" and in_first:
                if tok[1] == "list":
                    raise ValueError(f"{path}: list properties in the vertex element are not supported")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt == "binary_little_endian":
            rec = np.dtype(props)
            data = np.frombuffer(f.read(rec.itemsize * count), dtype=rec, count=count)
        elif fmt == "ascii":
            rows = np.loadtxt(f, max_rows=count, ndmin=2)
            data = {n: rows[:, i] for i, (n, _) in enumerate(props)}
        else:
            raise ValueError(f"{path}: unsupported PLY format {fmt}")
    return {n: np.asarray(data[n]) for n, _ in props}


def load_ply(path: str, max_sh_degree: int = 3, use_isotropic_gs: bool = False, fea_dim: int = 0, device="cpu") -> Dict[str, torch.Tensor]:
    """GaussianModel.load_ply (:283-336): returns the raw tensors ``xyz (P,3), features_dc (P,1,3), features_rest
    (P,M-1,3), opacity (P,1), scaling (P,3|1), rotation (P,4)`` [, ``feature (P,fea_dim)``] as float32 on ``device``."""
    v = read_ply_vertices(path)
    P = v["x"].shape[0]
    xyz = np.stack([v["x"], v["y"], v["z"]], 1)
    dc = np.stack([v["f_dc_0"], v["f_dc_1"], v["f_dc_2"]], 1)[:, :, None]                      # (P,3,1)
    # selected by prefix, in file order, like the reference's comprehension over the properties
    rest_names = [n for n in v if n.startswith("f_rest_")]
    if len(rest_names) != 3 * (max_sh_degree + 1) ** 2 - 3:
        raise ValueError(f"{path}: {len(rest_names)} f_rest_* properties do not fit SH degree {max_sh_degree}")
    rest = np.stack([v[n] for n in rest_names], 1).reshape(P, 3, (max_sh_degree + 1) ** 2 - 1) if rest_names else np.zeros((P, 3, 0))
    scales = np.stack([v[n] for n in v if n.startswith("scale_")], 1)
    if use_isotropic_gs:
        scales = scales[:, :1]
    rots = np.stack([v[n] for n in v if n.startswith("rot")], 1)
    t = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=torch.float32, device=device)
    out = {"xyz": t(xyz), "features_dc": t(dc).transpose(1, 2).contiguous(), "features_rest": t(rest).transpose(1, 2).contiguous(),
           "opacity": t(v["opacity"][:, None]), "scaling": t(scales), "rotation": t(rots)}
    if fea_dim > 0:
        fea = np.zeros((P, fea_dim))
        for i, n in enumerate([n for n in v if n.startswith("fea")]):
            fea[:, i] = v[n]
        out["feature"] = t(fea)
    return out


def save_skeleton_tree(path: str, joints, parent_indices, joint_node_indices=None, template_idx: int = 0) -> None:
    """``skeleton_tree.npz`` as train_rig.py:233 writes it: nodes (J,3), parents (J,), indices (J,), template_idx."""
    to_np = lambda t: t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)
    J = to_np(joints).shape[0]
    idx = to_np(joint_node_indices) if joint_node_indices is not None else np.arange(J)
    np.savez(path, nodes=to_np(joints), parents=to_np(parent_indices), indices=idx, template_idx=int(template_idx))


def load_skeleton_tree(path: str) -> Dict[str, object]:
    """Reads ``skeleton_tree.npz`` (train_rig.py:245-250; also the ``parent_indices`` key of the extraction step's
    file, skeleton_utils/mst_utils.py:348)."""
    z = np.load(path)
    parents = z["parent_indices"] if "parent_indices" in z.files else z["parents"]
    out = {"joints": torch.tensor(z["nodes"], dtype=torch.float32), "parents": torch.tensor(parents).long()}
    if "indices" in z.files:
        out["indices"] = torch.tensor(z["indices"]).long()
    if "template_idx" in z.files:
        out["template_idx"] = int(z["template_idx"])
    return out


def skeleton_state_dict(warp) -> Dict[str, torch.Tensor]:
    """State dict of a ``riggs_b200.skeleton_warp.SkeletonWarp`` under the reference's key names (``nodes``,
    ``_node_radius`` and the sub-modules' parameters), as SkeletonModel.save_weights stores it in ``skeleton.pth``."""
    return {k: v.detach().cpu() for k, v in warp.state_dict().items() if k not in ("parents", "parents_i32")}


def load_skeleton_state_dict(warp, state: Dict[str, torch.Tensor], strict: bool = False):
    """Loads the reference's ``skeleton.pth`` (torch.load of it) into a SkeletonWarp: ``nodes`` and ``_node_radius`` by
    name, MLP weights where the module names coincide; keys that have no counterpart here are returned."""
    own = warp.state_dict()
    unknown = []
    with torch.no_grad():
        for k, v in state.items():
            if k in own and tuple(own[k].shape) == tuple(v.shape):
                own[k].copy_(v.to(own[k].device, own[k].dtype))
            else:
                unknown.append(k)
    if strict and unknown:
        raise KeyError(f"skeleton state dict: no counterpart for {unknown}")
    return unknown
