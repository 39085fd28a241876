"""Writes a scene in the on-disk layout the reference's loader reads, from synthetic cameras and frames:

    <root>/sparse/0/cameras.bin, images.bin, points3D.bin      COLMAP binary model, one PINHOLE camera
    <root>/images/NNNNN.png          intensity frames            (scene/dataset_readers.py:152-154)
    <root>/images_event/NNNNN.png    frames the event loss differences (--event, :160-164)
    <root>/images_blurry/NNNNN.png   blurry frames (--deblur, :155-159)
    <root>/renders/NNNNN.png         test frames, read unconditionally (:177-179)

so that `python train.py -s <root> --event --gray [--deblur]` of the reference runs on data made here (SURVEY.md
section 8f rank 4: there is no network for the published datasets). The COLMAP binary format is public
(colmap: src/base/reconstruction.cc; read back by scene/colmap_loader.py:125-243): little-endian, counts as uint64,
camera = {int32 id, int32 model, uint64 width, uint64 height, float64 params[]} with model 1 = PINHOLE (fx, fy, cx,
cy); image = {int32 id, float64 qvec[4] (w, x, y, z of the world-to-camera rotation), float64 tvec[3], int32
camera_id, name + NUL, uint64 n_points2D, ...}; point = {uint64 id, float64 xyz[3], uint8 rgb[3], float64 error,
uint64 track_length, ...}. Pure numpy + PIL; no GPU needed.
"""
import os
import struct

import numpy as np

FRAME_DIRS = ("images", "images_event", "images_blurry", "renders")


def rotmat2qvec(R):
    """Unit quaternion (w, x, y, z) of a rotation matrix, w >= 0 (inverse of colmap_loader.qvec2rotmat)."""
    R = np.asarray(R, np.float64)
    K = np.array([[R[0, 0] - R[1, 1] - R[2, 2], 0, 0, 0],
                  [R[1, 0] + R[0, 1], R[1, 1] - R[0, 0] - R[2, 2], 0, 0],
                  [R[2, 0] + R[0, 2], R[2, 1] + R[1, 2], R[2, 2] - R[0, 0] - R[1, 1], 0],
                  [R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1], R[0, 0] + R[1, 1] + R[2, 2]]]) / 3.0
    w, v = np.linalg.eigh(K)
    q = v[[3, 0, 1, 2], np.argmax(w)]
    return -q if q[0] < 0 else q


def write_cameras_bin(path, width, height, fx, fy, cx, cy, camera_id=1):
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", 1))
        f.write(struct.pack("<iiQQ", camera_id, 1, width, height))          # model 1 = PINHOLE
        f.write(struct.pack("<dddd", fx, fy, cx, cy))


def write_images_bin(path, poses, names, camera_id=1):
    """poses: list of (R_w2c [3,3], t_w2c [3])."""
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(poses)))
        for i, ((R, t), name) in enumerate(zip(poses, names)):
            q = rotmat2qvec(R)
            f.write(struct.pack("<idddddddi", i + 1, q[0], q[1], q[2], q[3], float(t[0]), float(t[1]), float(t[2]), camera_id))
            f.write(name.encode("utf-8") + b"\x00")
            f.write(struct.pack("<Q", 0))                                     # no 2-D observations


def write_points3d_bin(path, xyz, rgb_u8):
    xyz = np.asarray(xyz, np.float64)
    rgb_u8 = np.asarray(rgb_u8, np.uint8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", xyz.shape[0]))
        for i in range(xyz.shape[0]):
            f.write(struct.pack("<QdddBBBd", i + 1, xyz[i, 0], xyz[i, 1], xyz[i, 2], int(rgb_u8[i, 0]), int(rgb_u8[i, 1]),
                                int(rgb_u8[i, 2]), 0.0))
            f.write(struct.pack("<Q", 0))                                     # empty track


def write_scene(root, cameras, frames_for, points, colors, dirs=FRAME_DIRS):
    """cameras: the dicts of synthetic.make_camera (R = camera-to-world rotation, T = world-to-camera translation,
    FoVx, FoVy, image_width, image_height: the reference's Camera convention, scene/cameras.py:17-57).
    frames_for(i) -> {dir name: uint8 [H,W,3] or float [3,H,W] in [0,1]} for view i (missing names reuse "images").
    points [P,3], colors [P,3] in [0,1]: the sparse cloud create_from_pcd starts from. Returns the image names."""
    from PIL import Image
    W, H = int(cameras[0]["image_width"]), int(cameras[0]["image_height"])
    sparse = os.path.join(root, "sparse", "0")
    os.makedirs(sparse, exist_ok=True)
    for d in dirs:
        os.makedirs(os.path.join(root, d), exist_ok=True)
    fx = W / (2.0 * np.tan(cameras[0]["FoVx"] * 0.5))
    fy = H / (2.0 * np.tan(cameras[0]["FoVy"] * 0.5))
    write_cameras_bin(os.path.join(sparse, "cameras.bin"), W, H, fx, fy, W / 2.0, H / 2.0)
    names = ["%05d.png" % i for i in range(len(cameras))]
    poses = [(np.asarray(c["R"], np.float64).T, np.asarray(c["T"], np.float64)) for c in cameras]
    write_images_bin(os.path.join(sparse, "images.bin"), poses, names)
    write_points3d_bin(os.path.join(sparse, "points3D.bin"), points, np.clip(np.round(np.asarray(colors) * 255.0), 0, 255))

    def to_u8(a):
        a = np.asarray(a)
        if a.dtype != np.uint8:
            if a.ndim == 3 and a.shape[0] == 3:
                a = np.transpose(a, (1, 2, 0))
            a = np.clip(np.round(a * 255.0), 0, 255).astype(np.uint8)
        return np.ascontiguousarray(a)

    for i, name in enumerate(names):
        fr = frames_for(i)
        for d in dirs:
            Image.fromarray(to_u8(fr.get(d, fr["images"]))).save(os.path.join(root, d, name))
    return names
