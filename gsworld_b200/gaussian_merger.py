"""Scene loading: 3DGS PLY files + semantic labels -> one merged model.

Mirrors the reference's loader for the observation path:
  * `GaussianModelMerger.load_models_from_config / merge_models`   gsworld/utils/gaussian_merger.py:155-274
    (config JSON = {"models": [{"data_path", "semantic_labels", "transformation"}, ...]}, e.g.
    configs/fr3_pnp_box.json; labels are an `.npy` path, an int, or absent; the "transformation" key is
    read but never applied by the reference, :176-177, and is ignored here too)
  * `Semantic3DGSWrapper.load_ply / save_ply`                       gsworld/mani_skill/utils/wrappers/semantic_3dgs_wrapper.py:60-167
    (PLY vertex properties x,y,z,nx,ny,nz,f_dc_0..2,f_rest_0..44,opacity,scale_0..2,rot_0..3[,semantics], all
    float32; f_rest is channel-major in the file; tensor layouts of SURVEY.md Appendix A).
The reference goes through the `plyfile` package and per-property Python loops; here the binary
little-endian vertex block is mapped with one numpy structured dtype.  Not on the per-step path.
"""
from __future__ import annotations

import json
import os
from typing import List, Optional

import numpy as np
import torch

from .fixtures import GaussianScene

_PLY_TYPES = {"float": "<f4", "float32": "<f4", "double": "<f8", "float64": "<f8", "uchar": "u1", "uint8": "u1",
              "char": "i1", "int8": "i1", "short": "<i2", "int16": "<i2", "ushort": "<u2", "uint16": "<u2",
              "int": "<i4", "int32": "<i4", "uint": "<u4", "uint32": "<u4"}


def read_ply_vertices(path: str) -> np.ndarray:
    """Vertex element of a binary_little_endian (or ascii) PLY as a numpy structured array."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, count, props, in_vertex = None, None, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: truncated PLY header")
            tok = line.decode("ascii", "replace").split()
            if not tok:
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_vertex = tok[1] == "vertex"
                if in_vertex:
                    count = int(tok[2])
                elif count is None:
                    raise ValueError(f"{path}: an element precedes 'vertex'; unsupported layout")
            elif tok[0] == "property" and in_vertex:
                if tok[1] == "list":
                    raise ValueError(f"{path}: list property in the vertex element")
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if count is None:
            raise ValueError(f"{path}: no vertex element")
        dtype = np.dtype(props)
        if fmt == "binary_little_endian":
            data = np.fromfile(f, dtype=dtype, count=count)
        elif fmt == "ascii":
            data = np.loadtxt(f, dtype=np.float64, max_rows=count, ndmin=2)
            out = np.empty(count, dtype=dtype)
            for k, (n, _) in enumerate(props):
                out[n] = data[:, k]
            data = out
        else:
            raise ValueError(f"{path}: unsupported PLY format {fmt}")
        if data.shape[0] != count:
            raise ValueError(f"{path}: expected {count} vertices, file holds {data.shape[0]}")
        return data


def load_ply(path: str, device="cuda", max_sh_degree: int = 3) -> GaussianScene:
    """Semantic3DGSWrapper.load_ply: returns the model with `_semantics` (N,1) float (zeros if absent)."""
    v = read_ply_vertices(path)
    names = v.dtype.names
    n = v.shape[0]
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    xyz = np.stack([v["x"], v["y"], v["z"]], 1)
    rest_names = sorted([p for p in names if p.startswith("f_rest_")], key=lambda s: int(s.split("_")[-1]))
    n_rest = 3 * (max_sh_degree + 1) ** 2 - 3
    if len(rest_names) != n_rest:
        raise AssertionError(f"{path}: {len(rest_names)} f_rest properties, expected {n_rest}")
    dc = np.stack([v["f_dc_0"], v["f_dc_1"], v["f_dc_2"]], 1)[:, None, :]                   # (N,1,3)
    rest = np.stack([v[p] for p in rest_names], 1).reshape(n, 3, n_rest // 3).transpose(0, 2, 1)  # (N,15,3)
    scale_names = sorted([p for p in names if p.startswith("scale_")], key=lambda s: int(s.split("_")[-1]))
    rot_names = sorted([p for p in names if p.startswith("rot")], key=lambda s: int(s.split("_")[-1]))
    scales = np.stack([v[p] for p in scale_names], 1)
    rots = np.stack([v[p] for p in rot_names], 1)
    opac = np.asarray(v["opacity"])[:, None, None]                                           # (N,1,1)
    sem = np.asarray(v["semantics"])[:, None] if "semantics" in names else np.zeros((n, 1))
    t = lambda a: torch.from_numpy(f32(a)).to(device)
    return GaussianScene(t(xyz), t(dc), t(rest), t(opac), t(scales), t(rots), t(sem))


def save_ply(model: GaussianScene, path: str, with_semantics: bool = True) -> None:
    """Semantic3DGSWrapper.save_ply layout (semantic_3dgs_wrapper.py:60-98): binary little endian, float32."""
    n = model._xyz.shape[0]
    c = lambda t: t.detach().float().cpu().numpy()
    dc = c(model._features_dc).transpose(0, 2, 1).reshape(n, -1)        # (N,3)
    rest = c(model._features_rest).transpose(0, 2, 1).reshape(n, -1)    # channel-major (N,45)
    cols = [("x", c(model._xyz)[:, 0]), ("y", c(model._xyz)[:, 1]), ("z", c(model._xyz)[:, 2]),
            ("nx", np.zeros(n)), ("ny", np.zeros(n)), ("nz", np.zeros(n))]
    cols += [(f"f_dc_{k}", dc[:, k]) for k in range(dc.shape[1])]
    cols += [(f"f_rest_{k}", rest[:, k]) for k in range(rest.shape[1])]
    cols += [("opacity", c(model._opacity).reshape(n))]
    cols += [(f"scale_{k}", c(model._scaling)[:, k]) for k in range(3)]
    cols += [(f"rot_{k}", c(model._rotation)[:, k]) for k in range(4)]
    if with_semantics:
        cols += [("semantics", c(model._semantics).reshape(n))]
    arr = np.empty(n, dtype=[(k, "<f4") for k, _ in cols])
    for k, a in cols:
        arr[k] = a
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(b"ply\nformat binary_little_endian 1.0\n")
        f.write(f"element vertex {n}\n".encode())
        for k, _ in cols:
            f.write(f"property float {k}\n".encode())
        f.write(b"end_header\n")
        arr.tofile(f)


class GaussianModelMerger:
    """Same surface as the reference class (gaussian_merger.py:13-323) for the calls GSWorldWrapper makes."""

    def __init__(self, device="cuda", asset_dir: Optional[str] = None):
        self.device = device
        self.asset_dir = asset_dir or os.environ.get("GSWORLD_ASSET_DIR", ".")
        self.models: List[GaussianScene] = []
        self.model_paths: List[str] = []
        self.model_configs: List[dict] = []
        self.merged_model: Optional[GaussianScene] = None

    def load_config_from_json(self, json_path):
        if not os.path.exists(json_path):
            raise FileNotFoundError(f"JSON configuration file not found: {json_path}")
        try:
            with open(json_path, "r") as f:
                config = json.load(f)
        except json.JSONDecodeError:
            raise ValueError(f"Invalid JSON format in file: {json_path}")
        if "models" not in config or not isinstance(config["models"], list):
            raise ValueError("JSON file should contain a 'models' list")
        self.model_configs = list(config["models"])
        return self.model_configs

    def assign_semantic_labels(self, model, semantic_labels):
        n = model._xyz.shape[0]
        if isinstance(semantic_labels, str):
            assert os.path.exists(semantic_labels)
            labels = np.load(semantic_labels)
            model._semantics = torch.from_numpy(labels).to(self.device)[..., None]
        elif isinstance(semantic_labels, (int, float)):
            model._semantics = torch.full((n, 1), float(semantic_labels), device=self.device)
        elif isinstance(semantic_labels, torch.Tensor):
            model._semantics = semantic_labels.to(self.device)
        else:
            model._semantics = torch.zeros(n, 1, device=self.device)
        return model

    def load_model_from_config(self, model_config):
        if "data_path" not in model_config:
            raise ValueError("Missing required 'data_path' in model config")
        ply_path = os.path.join(self.asset_dir, model_config["data_path"])
        if not os.path.exists(ply_path):
            raise FileNotFoundError(f"PLY file not found: {ply_path}")
        semantic_labels = model_config.get("semantic_labels", None)
        model = load_ply(ply_path, device=self.device)
        if semantic_labels is not None:
            if isinstance(semantic_labels, str):
                semantic_labels = os.path.join(self.asset_dir, semantic_labels)
            model = self.assign_semantic_labels(model, semantic_labels)
        self.models.append(model)
        self.model_paths.append(ply_path)
        return len(self.models) - 1

    def load_multiple_models(self, model_configs):
        return [self.load_model_from_config(c) for c in model_configs]

    def load_models_from_config(self, json_path):
        return self.load_multiple_models(self.load_config_from_json(json_path))

    def get_model(self, index):
        if 0 <= index < len(self.models):
            return self.models[index]
        raise IndexError(f"Model index {index} is out of range")

    def merge_models(self, indices=None):
        if not self.models:
            raise ValueError("No models to merge")
        ms = self.models if indices is None else [self.get_model(i) for i in indices]
        if not ms:
            raise ValueError("No valid models to merge")
        cat = lambda name: torch.cat([getattr(m, name) for m in ms], dim=0)
        # the reference concatenates int64 npy labels with float constants; torch.cat promotes to float
        sem = torch.cat([m._semantics.to(torch.float32) if m._semantics.dtype != torch.float64 else m._semantics
                         for m in ms], dim=0)
        self.merged_model = GaussianScene(cat("_xyz"), cat("_features_dc"), cat("_features_rest"), cat("_opacity"),
                                          cat("_scaling"), cat("_rotation"), sem)
        return self.merged_model

    def get_merged_model(self):
        if self.merged_model is None:
            raise ValueError("No merged model exists. Call merge_models() first.")
        return self.merged_model

    def save_merged_model(self, output_path):
        save_ply(self.get_merged_model(), output_path)
        return True

    def clear_models(self):
        self.models, self.model_paths, self.model_configs, self.merged_model = [], [], [], None
