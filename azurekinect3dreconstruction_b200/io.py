"""o3d.io look-alikes and the reference's own on-disk formats.  Host-side only: results are
read back from the device once and written with numpy; nothing here touches the hot path.

Open3D-shaped calls the reference makes (37 write_point_cloud, 27 write_triangle_mesh sites; `.ply` and `.obj` only):
    o3d.io.write_point_cloud(path, pcd)          CodeThatWorks/GoodColorReconstruction.py:543,563
    o3d.io.write_triangle_mesh(path, mesh)       GoodColorReconstruction.py:538   (OBJ), CreateMesh.py (PLY)
    o3d.io.read_point_cloud / read_triangle_mesh LastReconstructionVisualizer.py, ChooseReconstruction.py
Formats follow what Open3D 0.19 writes so its viewers can load the files on a machine that has a display:
  * PLY: `format binary_little_endian 1.0` (or ascii), `comment Created by Open3D`, vertex properties `double x y z`,
    then `double nx ny nz` when normals exist, then `uchar red green blue` when colours exist; meshes add
    `element face N` / `property list uchar uint vertex_indices`;
  * OBJ: `v x y z [r g b]`, `vn nx ny nz`, faces `f a//a b//b c//c` (1-based) when normals exist else `f a b c`.
The reference's own files:
  * raw frames   `np.savez(path, color=BGRA u8 HxWx4, depth=u16 HxW)`           CodeThatKindaWorks/Depth2PC.py:190-192
  * trajectory   index line + four rows of the 4x4 pose per frame               GoodColorReconstruction.py:549-556
  * calibration  JSON {timestamp, num_cameras, camera_serials, intrinsics[], extrinsics[]}
                                                                                InitialReconstruction-CUDA-2cam:544-565
"""
from __future__ import annotations

import json
import os
import time
from typing import List, Optional, Sequence

import numpy as np

from .camera import PinholeCameraIntrinsic
from .geometry import PointCloud, TriangleMesh
from .utility import Vector3dVector, Vector3iVector


def _colors_u8(c: np.ndarray) -> np.ndarray:
    # Open3D: (uint8_t)std::max(std::min((int)(color * 255.0), 255), 0)  -- truncation, not rounding
    return np.clip((np.asarray(c, dtype=np.float64) * 255.0).astype(np.int64), 0, 255).astype(np.uint8)


def _ply_header(n_vertex: int, has_normals: bool, has_colors: bool, n_face: Optional[int], ascii_: bool) -> bytes:
    lines = ["ply", "format %s 1.0" % ("ascii" if ascii_ else "binary_little_endian"), "comment Created by Open3D",
             "element vertex %d" % n_vertex, "property double x", "property double y", "property double z"]
    if has_normals:
        lines += ["property double nx", "property double ny", "property double nz"]
    if has_colors:
        lines += ["property uchar red", "property uchar green", "property uchar blue"]
    if n_face is not None:
        lines += ["element face %d" % n_face, "property list uchar uint vertex_indices"]
    lines.append("end_header")
    return ("\n".join(lines) + "\n").encode("ascii")


def _vertex_records(xyz, nrm, col) -> np.ndarray:
    fields = [("x", "<f8"), ("y", "<f8"), ("z", "<f8")]
    if nrm is not None:
        fields += [("nx", "<f8"), ("ny", "<f8"), ("nz", "<f8")]
    if col is not None:
        fields += [("red", "u1"), ("green", "u1"), ("blue", "u1")]
    rec = np.empty(len(xyz), dtype=np.dtype(fields))
    rec["x"], rec["y"], rec["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    if nrm is not None:
        rec["nx"], rec["ny"], rec["nz"] = nrm[:, 0], nrm[:, 1], nrm[:, 2]
    if col is not None:
        u8 = _colors_u8(col)
        rec["red"], rec["green"], rec["blue"] = u8[:, 0], u8[:, 1], u8[:, 2]
    return rec


def _write_ply(path, xyz, nrm, col, tri, ascii_):
    with open(path, "wb") as f:
        f.write(_ply_header(len(xyz), nrm is not None, col is not None, None if tri is None else len(tri), ascii_))
        if ascii_:
            u8 = _colors_u8(col) if col is not None else None
            for i in range(len(xyz)):
                parts = ["%.10f" % v for v in xyz[i]]          # Open3D prints doubles with 10 decimals in ascii PLY
                if nrm is not None:
                    parts += ["%.10f" % v for v in nrm[i]]
                if u8 is not None:
                    parts += [str(int(v)) for v in u8[i]]
                f.write((" ".join(parts) + "\n").encode("ascii"))
            if tri is not None:
                for t in tri:
                    f.write(("3 %d %d %d\n" % (t[0], t[1], t[2])).encode("ascii"))
        else:
            f.write(_vertex_records(xyz, nrm, col).tobytes())
            if tri is not None:
                rec = np.empty(len(tri), dtype=np.dtype([("n", "u1"), ("v", "<u4", (3,))]))
                rec["n"] = 3
                rec["v"] = tri.astype(np.uint32)
                f.write(rec.tobytes())


_PLY_TYPES = {"char": "i1", "uchar": "u1", "short": "<i2", "ushort": "<u2", "int": "<i4", "uint": "<u4", "float": "<f4", "double": "<f8",
              "int8": "i1", "uint8": "u1", "int16": "<i2", "uint16": "<u2", "int32": "<i4", "uint32": "<u4", "float32": "<f4", "float64": "<f8"}


def _read_ply(path):
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise RuntimeError("Read PLY failed: unable to open file: %s" % path)
        fmt, elements = None, []
        while True:
            line = f.readline()
            if not line:
                raise RuntimeError("Read PLY failed: header is not terminated")
            tok = line.decode("ascii", "replace").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                elements.append([tok[1], int(tok[2]), []])
            elif tok[0] == "property":
                elements[-1][2].append(tok[1:])
            elif tok[0] == "end_header":
                break
        if fmt not in ("ascii", "binary_little_endian"):
            raise RuntimeError("Read PLY failed: unsupported format %s" % fmt)
        out = {}
        for name, count, props in elements:
            is_list = any(p[0] == "list" for p in props)
            if not is_list:
                dt = np.dtype([(p[1], _PLY_TYPES[p[0]]) for p in props])
                if fmt == "ascii":
                    rows = [f.readline().split() for _ in range(count)]
                    rec = np.empty(count, dtype=dt)
                    for j, p in enumerate(props):
                        rec[p[1]] = [r[j] for r in rows] if count else []
                else:
                    rec = np.frombuffer(f.read(dt.itemsize * count), dtype=dt, count=count)
                out[name] = rec
            else:
                p = props[0]                                     # property list <count type> <index type> vertex_indices
                if fmt == "ascii":
                    faces = [[int(v) for v in f.readline().split()[1:4]] for _ in range(count)]
                    out[name] = np.asarray(faces, dtype=np.int32).reshape(-1, 3)
                else:
                    dt = np.dtype([("n", _PLY_TYPES[p[1]]), ("v", _PLY_TYPES[p[2]], (3,))])
                    rec = np.frombuffer(f.read(dt.itemsize * count), dtype=dt, count=count)
                    if count and not (rec["n"] == 3).all():
                        raise RuntimeError("Read PLY failed: only triangle faces are supported")
                    out[name] = rec["v"].astype(np.int32)
        return out


def _vertex_arrays(rec):
    names = rec.dtype.names or ()
    xyz = np.stack([rec["x"], rec["y"], rec["z"]], 1).astype(np.float64) if len(rec) else np.zeros((0, 3))
    nrm = np.stack([rec["nx"], rec["ny"], rec["nz"]], 1).astype(np.float64) if "nx" in names else None
    col = np.stack([rec["red"], rec["green"], rec["blue"]], 1).astype(np.float64) / 255.0 if "red" in names else None
    return xyz, nrm, col


def _ext(path: str) -> str:
    return os.path.splitext(path)[1].lower()


# ---- point clouds ---------------------------------------------------------------------------------------------
def write_point_cloud(filename: str, pointcloud: PointCloud, write_ascii: bool = False, compressed: bool = False,
                      print_progress: bool = False) -> bool:
    """o3d.io.write_point_cloud (PLY).  Returns True on success, False (with a warning) on an unknown extension, as upstream."""
    if _ext(filename) != ".ply":
        print("[Open3D WARNING] Write geometry::PointCloud failed: unknown file extension %s." % _ext(filename))
        return False
    xyz = np.asarray(pointcloud.points, dtype=np.float64).reshape(-1, 3)
    nrm = np.asarray(pointcloud.normals, dtype=np.float64) if pointcloud.has_normals() else None
    col = np.asarray(pointcloud.colors, dtype=np.float64) if pointcloud.has_colors() else None
    _write_ply(filename, xyz, nrm, col, None, write_ascii)
    return True


def read_point_cloud(filename: str, format: str = "auto", remove_nan_points: bool = False, remove_infinite_points: bool = False,
                     print_progress: bool = False) -> PointCloud:
    pcd = PointCloud()
    if _ext(filename) != ".ply" or not os.path.exists(filename):
        print("[Open3D WARNING] Read geometry::PointCloud failed: unknown file extension or missing file: %s" % filename)
        return pcd
    xyz, nrm, col = _vertex_arrays(_read_ply(filename)["vertex"])
    keep = np.ones(len(xyz), bool)
    if remove_nan_points:
        keep &= ~np.isnan(xyz).any(1)
    if remove_infinite_points:
        keep &= ~np.isinf(xyz).any(1)
    pcd.points = Vector3dVector(xyz[keep])
    if nrm is not None:
        pcd.normals = Vector3dVector(nrm[keep])
    if col is not None:
        pcd.colors = Vector3dVector(col[keep])
    return pcd


# ---- triangle meshes --------------------------------------------------------------------------------------------
def write_triangle_mesh(filename: str, mesh: TriangleMesh, write_ascii: bool = False, compressed: bool = False,
                        write_vertex_normals: bool = True, write_vertex_colors: bool = True, write_triangle_uvs: bool = True,
                        print_progress: bool = False) -> bool:
    """o3d.io.write_triangle_mesh (.ply, .obj)."""
    ext = _ext(filename)
    xyz = np.asarray(mesh.vertices, dtype=np.float64).reshape(-1, 3)
    tri = np.asarray(mesh.triangles, dtype=np.int64).reshape(-1, 3)
    nrm = np.asarray(mesh.vertex_normals, dtype=np.float64) if (write_vertex_normals and mesh.has_vertex_normals()) else None
    col = np.asarray(mesh.vertex_colors, dtype=np.float64) if (write_vertex_colors and mesh.has_vertex_colors()) else None
    if ext == ".ply":
        _write_ply(filename, xyz, nrm, col, tri, write_ascii)
        return True
    if ext == ".obj":
        with open(filename, "w") as f:
            f.write("# Created by Open3D \n# object name: %s\n" % os.path.splitext(os.path.basename(filename))[0])
            f.write("# number of vertices: %d\n# number of triangles: %d\n" % (len(xyz), len(tri)))
            chunks = []
            for i in range(len(xyz)):
                if col is not None:
                    chunks.append("v %.10g %.10g %.10g %.10g %.10g %.10g\n" % (*xyz[i], *col[i]))
                else:
                    chunks.append("v %.10g %.10g %.10g\n" % tuple(xyz[i]))
                if nrm is not None:
                    chunks.append("vn %.10g %.10g %.10g\n" % tuple(nrm[i]))
            f.write("".join(chunks))
            t1 = tri + 1
            if nrm is not None:
                f.write("".join("f %d//%d %d//%d %d//%d\n" % (a, a, b, b, c, c) for a, b, c in t1))
            else:
                f.write("".join("f %d %d %d\n" % (a, b, c) for a, b, c in t1))
        return True
    print("[Open3D WARNING] Write geometry::TriangleMesh failed: unknown file extension %s." % ext)
    return False


def read_triangle_mesh(filename: str, enable_post_processing: bool = False, print_progress: bool = False) -> TriangleMesh:
    mesh = TriangleMesh()
    ext = _ext(filename)
    if not os.path.exists(filename) or ext not in (".ply", ".obj"):
        print("[Open3D WARNING] Read geometry::TriangleMesh failed: unknown file extension or missing file: %s" % filename)
        return mesh
    if ext == ".ply":
        data = _read_ply(filename)
        xyz, nrm, col = _vertex_arrays(data["vertex"])
        tri = data.get("face", np.zeros((0, 3), np.int32))
    else:
        v, vn, vc, faces = [], [], [], []
        with open(filename) as f:
            for line in f:
                t = line.split()
                if not t:
                    continue
                if t[0] == "v":
                    v.append([float(x) for x in t[1:4]])
                    if len(t) >= 7:
                        vc.append([float(x) for x in t[4:7]])
                elif t[0] == "vn":
                    vn.append([float(x) for x in t[1:4]])
                elif t[0] == "f":
                    faces.append([int(x.split("/")[0]) - 1 for x in t[1:4]])
        xyz = np.asarray(v, np.float64).reshape(-1, 3)
        nrm = np.asarray(vn, np.float64) if len(vn) == len(v) and v else None
        col = np.asarray(vc, np.float64) if len(vc) == len(v) and v else None
        tri = np.asarray(faces, np.int32).reshape(-1, 3)
    mesh.vertices = Vector3dVector(xyz)
    mesh.triangles = Vector3iVector(tri)
    if nrm is not None:
        mesh.vertex_normals = Vector3dVector(nrm)
    if col is not None:
        mesh.vertex_colors = Vector3dVector(col)
    return mesh


# ---- the reference's own files ---------------------------------------------------------------------------------
def save_frame_npz(path: str, color_bgra: np.ndarray, depth_u16: np.ndarray) -> None:
    """Depth2PC.py:190-192: np.savez(frame_path, color=capture.color, depth=capture.transformed_depth)."""
    np.savez(path, color=np.asarray(color_bgra, dtype=np.uint8), depth=np.asarray(depth_u16, dtype=np.uint16))


def load_frame_npz(path: str):
    """Depth2PC.py:224-240 replay: returns (color BGRA u8, depth u16)."""
    with np.load(path) as z:
        return z["color"], z["depth"]


def write_trajectory(path: str, poses: Sequence[np.ndarray]) -> None:
    """GoodColorReconstruction.py:549-556: an index line, then the four rows of the pose, `str()` of each float."""
    with open(path, "w") as f:
        for i, pose in enumerate(poses):
            pose = np.asarray(pose, dtype=np.float64)
            f.write("%d\n" % i)
            for j in range(4):
                f.write(" ".join(str(pose[j, k]) for k in range(4)) + "\n")


def read_trajectory(path: str) -> List[np.ndarray]:
    with open(path) as f:
        lines = [ln for ln in (x.strip() for x in f) if ln]
    if len(lines) % 5:
        raise RuntimeError("trajectory file %s is truncated" % path)
    return [np.array([[float(v) for v in lines[i + 1 + j].split()] for j in range(4)], dtype=np.float64) for i in range(0, len(lines), 5)]


def save_calibration(path: str, intrinsics: Sequence[PinholeCameraIntrinsic], extrinsics: Sequence[np.ndarray],
                     camera_serials: Optional[Sequence[str]] = None, timestamp: Optional[str] = None) -> dict:
    """InitialReconstruction-CUDA-2cam:544-575 schema."""
    data = {"timestamp": timestamp or time.strftime("%Y%m%d-%H%M%S"), "num_cameras": len(intrinsics),
            "camera_serials": list(camera_serials) if camera_serials is not None else ["synthetic-%d" % i for i in range(len(intrinsics))],
            "intrinsics": [], "extrinsics": []}
    for K in intrinsics:
        fx, fy = K.get_focal_length()
        cx, cy = K.get_principal_point()
        data["intrinsics"].append({"width": K.width, "height": K.height, "fx": fx, "fy": fy, "cx": cx, "cy": cy})
    for E in extrinsics:
        data["extrinsics"].append(np.asarray(E, dtype=np.float64).tolist())
    with open(path, "w") as f:
        json.dump(data, f, indent=4)
    return data


def load_calibration(path: str):
    """Returns (intrinsics, extrinsics, raw dict)  (InitialReconstruction-CUDA-2cam:577-623)."""
    with open(path) as f:
        data = json.load(f)
    Ks = [PinholeCameraIntrinsic(d["width"], d["height"], d["fx"], d["fy"], d["cx"], d["cy"]) for d in data["intrinsics"]]
    Es = [np.asarray(e, dtype=np.float64) for e in data["extrinsics"]]
    return Ks, Es, data


# ---- pose-graph JSON (o3d.io.write_pose_graph / read_pose_graph; capture_frames.py:62-64, bundlefusion_reconstruction.py:284-286)
# Open3D's IJsonConvertible layout: matrices as flat arrays in Eigen's column-major coefficient order, every object
# tagged with class_name / version_major / version_minor.
def _pg_mat(a) -> list:
    return [float(v) for v in np.asarray(a, dtype=np.float64).reshape(a.shape).T.reshape(-1)]


def write_pose_graph(filename: str, pose_graph) -> bool:
    doc = {
        "class_name": "PoseGraph",
        "edges": [{"class_name": "PoseGraphEdge", "confidence": float(e.confidence), "information": _pg_mat(e.information),
                   "source_node_id": int(e.source_node_id), "target_node_id": int(e.target_node_id),
                   "transformation": _pg_mat(e.transformation), "uncertain": bool(e.uncertain),
                   "version_major": 1, "version_minor": 0} for e in pose_graph.edges],
        "nodes": [{"class_name": "PoseGraphNode", "pose": _pg_mat(nd.pose), "version_major": 1, "version_minor": 0}
                  for nd in pose_graph.nodes],
        "version_major": 1, "version_minor": 0,
    }
    with open(filename, "w") as f:
        json.dump(doc, f, indent="\t", sort_keys=True)
        f.write("\n")
    return True


def read_pose_graph(filename: str):
    from .pipelines.registration import PoseGraph, PoseGraphEdge, PoseGraphNode
    with open(filename) as f:
        doc = json.load(f)
    if doc.get("class_name") != "PoseGraph":
        raise RuntimeError("read_pose_graph: not a PoseGraph file: " + filename)
    pg = PoseGraph()
    for nd in doc.get("nodes", []):
        pg.nodes.append(PoseGraphNode(np.asarray(nd["pose"], dtype=np.float64).reshape(4, 4).T))
    for e in doc.get("edges", []):
        pg.edges.append(PoseGraphEdge(e["source_node_id"], e["target_node_id"], np.asarray(e["transformation"], dtype=np.float64).reshape(4, 4).T,
                                      np.asarray(e["information"], dtype=np.float64).reshape(6, 6).T, bool(e["uncertain"]), float(e["confidence"])))
    return pg


# ---------------------------------------------------------------- images (o3d.io.read_image / write_image)
def read_image(filename: str):
    """`o3d.io.read_image(path)` (CodeThatWorks/RGBD_Odometry.py:10-13): PNG/JPG -> geometry.Image.  8-bit colour
    comes back as RGB (Open3D's channel order), 16-bit single-channel depth PNGs stay u16.  A missing or
    undecodable file gives an empty Image plus a warning upstream; here it raises, since every caller goes
    straight on to create_from_color_and_depth, which would raise anyway."""
    import cv2
    from .geometry import Image
    a = cv2.imread(filename, cv2.IMREAD_UNCHANGED)
    if a is None:
        raise RuntimeError(f"read_image: cannot read {filename}")
    if a.ndim == 3:
        a = a[:, :, 2::-1] if a.shape[2] >= 3 else a[:, :, 0]  # BGR(A) -> RGB, alpha dropped like upstream's PNG reader
    return Image(np.ascontiguousarray(a))


def write_image(filename: str, image, quality: int = -1) -> bool:
    """`o3d.io.write_image(path, image)`: u8 (1 or 3 channels, RGB order) or u16 single-channel -> PNG/JPG."""
    import cv2
    a = image.numpy() if hasattr(image, "numpy") else np.asarray(image)
    if a.dtype not in (np.uint8, np.uint16):
        raise RuntimeError("write_image: only 8-bit and 16-bit integer images can be written")
    if a.ndim == 3 and a.shape[2] == 3:
        a = a[:, :, ::-1]
    params = [cv2.IMWRITE_JPEG_QUALITY, quality] if quality >= 0 and _ext(filename) in (".jpg", ".jpeg") else []
    return bool(cv2.imwrite(filename, np.ascontiguousarray(a), params))
