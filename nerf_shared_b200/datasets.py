"""Dataset loading (reference utils.load_datasets, utils.py:216-313, and its four disk loaders).

Host-side I/O, no device arithmetic: the on-disk formats of the reference are read into the same arrays
(`images, poses, render_poses, hwf, i_split, K, bds_dict`) so that every shipped `configs/*.txt` runs through the
unchanged main.py when its data directory is present.

* Blender / NeRF-synthetic: `transforms_{train,val,test}.json` + RGBA PNGs  (load_blender.py:44-98)
* LLFF: `poses_bounds.npy` + `images[_<factor>]/`                           (load_llff.py:62-319)
* LINEMOD: transforms JSON with per-frame intrinsics                         (load_LINEMOD.py:36-94)
* DeepVoxels: `intrinsics.txt`, `pose/*.txt`, `rgb/*.png`                    (load_deepvoxels.py:6-110)
* `synthetic`: no files, a Blender-shaped random scene (benchmarks and tests; not a reference type)

The arithmetic that fixes results (pose conventions, recentring, bound scaling, spiral / spherical render
paths, train / val / test splits, near / far) follows the reference line by line in meaning; images are
decoded with imageio when it is installed and with OpenCV otherwise (same bytes for PNG).
"""
from __future__ import annotations

import json
import os

import numpy as np


# ------------------------------------------------------------------------------------------ image files
def _imread(path):
    """-> uint8 array [H, W, C] in RGB(A) order."""
    try:
        import imageio
        if hasattr(imageio, "imread"):
            return np.asarray(imageio.imread(path))
    except Exception:
        pass
    import cv2
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img is None:
        raise FileNotFoundError(path)
    if img.ndim == 3 and img.shape[2] == 4:
        img = cv2.cvtColor(img, cv2.COLOR_BGRA2RGBA)
    elif img.ndim == 3:
        img = cv2.cvtColor(img, cv2.COLOR_BGR2RGB)
    return img


def _area_resize(img, W, H):
    import cv2
    return cv2.resize(img, (W, H), interpolation=cv2.INTER_AREA)


_IMG_EXT = ("JPG", "jpg", "png", "jpeg", "PNG")


# ------------------------------------------------------------------------------------------ camera paths
def _rot_x(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[1, 0, 0, 0], [0, c, -s, 0], [0, s, c, 0], [0, 0, 0, 1]], dtype=np.float32)


def _rot_y(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, -s, 0], [0, 1, 0, 0], [s, 0, c, 0], [0, 0, 0, 1]], dtype=np.float32)


def _pose_spherical(theta, phi, radius):
    """Orbit camera looking at the origin (load_LINEMOD.py:29-34): translate along z, pitch, yaw, axis swap."""
    t = np.eye(4, dtype=np.float32)
    t[2, 3] = radius
    c2w = _rot_y(np.float32(theta / 180.0 * np.pi)) @ (_rot_x(np.float32(phi / 180.0 * np.pi)) @ t)
    swap = np.array([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], dtype=np.float32)
    return (swap @ c2w).astype(np.float32)


def _pose_blender_path(theta, phi):
    """The render path load_blender.py:36-41 builds: pitch, yaw, then a fixed offset (3, 0.3, -1) -- no radius."""
    c2w = _rot_y(np.float32(theta / 180.0 * np.pi)) @ _rot_x(np.float32(phi / 180.0 * np.pi))
    off = np.array([[1, 0, 0, 3], [0, 1, 0, 0.3], [0, 0, 1, -1], [0, 0, 0, 1]], dtype=np.float32)
    return (off @ c2w).astype(np.float32)


# ------------------------------------------------------------------------------------------ Blender / LINEMOD
def _read_transforms(basedir, testskip, path_of):
    metas = {}
    for s in ("train", "val", "test"):
        with open(os.path.join(basedir, "transforms_%s.json" % s)) as f:
            metas[s] = json.load(f)
    imgs, poses, counts = [], [], [0]
    for s in ("train", "val", "test"):
        skip = 1 if (s == "train" or testskip == 0) else testskip
        frames = metas[s]["frames"][::skip]
        part = np.array([_imread(path_of(fr)) for fr in frames])
        imgs.append((part / 255.0).astype(np.float32))                     # all channels kept (RGBA)
        poses.append(np.array([fr["transform_matrix"] for fr in frames]).astype(np.float32))
        counts.append(counts[-1] + len(frames))
    i_split = [np.arange(counts[i], counts[i + 1]) for i in range(3)]
    return metas, np.concatenate(imgs, 0), np.concatenate(poses, 0), i_split


def _halve(imgs, H, W, channels):
    out = np.zeros((imgs.shape[0], H, W, channels))
    for i, img in enumerate(imgs):
        out[i] = _area_resize(img, W, H)
    return out


def load_blender_data(basedir, half_res=False, testskip=1):
    """-> imgs [N,H,W,4], poses [N,4,4], render_poses [40,4,4], [H,W,focal], i_split, near, far (load_blender.py:44-98).
    near / far come from the last split's JSON, as in the reference (:57)."""
    metas, imgs, poses, i_split = _read_transforms(
        basedir, testskip, lambda fr: os.path.join(basedir, fr["file_path"] + ".png"))
    last = metas["test"]
    near, far = last["near"], last["far"]
    H, W = imgs[0].shape[:2]
    focal = 0.5 * W / np.tan(0.5 * float(last["camera_angle_x"]))
    render_poses = np.stack([_pose_blender_path(a, 0.0) for a in np.linspace(-180, 180, 40 + 1)[:-1]], 0)
    if half_res:
        H, W, focal = H // 2, W // 2, focal / 2.0
        imgs = _halve(imgs, H, W, 4)
    return imgs, poses, render_poses, [H, W, focal], i_split, near, far


def load_LINEMOD_data(basedir, half_res=False, testskip=1):
    """-> imgs, poses, render_poses, [H,W,focal], K, i_split, near, far (load_LINEMOD.py:36-94); file paths in the
    JSON are used as written, K is the first test frame's intrinsic matrix."""
    metas, imgs, poses, i_split = _read_transforms(basedir, testskip, lambda fr: fr["file_path"])
    first = metas["test"]["frames"][0]
    H, W = imgs[0].shape[:2]
    K = first["intrinsic_matrix"]
    focal = float(K[0][0])
    render_poses = np.stack([_pose_spherical(a, -30.0, 4.0) for a in np.linspace(-180, 180, 40 + 1)[:-1]], 0)
    if half_res:
        H, W, focal = H // 2, W // 2, focal / 2.0
        imgs = _halve(imgs, H, W, 3)
    near = np.floor(min(metas["train"]["near"], metas["test"]["near"]))
    far = np.ceil(max(metas["train"]["far"], metas["test"]["far"]))
    return imgs, poses, render_poses, [H, W, focal], K, i_split, near, far


# ------------------------------------------------------------------------------------------ LLFF
def _unit(v):
    return v / np.linalg.norm(v)


def _look_at(z, up, pos):
    """[3,4] camera-to-world with third axis z, `up` as the approximate second axis (load_llff.py:122-128)."""
    z = _unit(z)
    x = _unit(np.cross(up, z))
    y = _unit(np.cross(z, x))
    return np.stack([x, y, z, pos], 1)


def _average_pose(poses):
    """[3,5]: mean centre, summed view / up axes, hwf column of the first pose (load_llff.py:134-144)."""
    centre = poses[:, :3, 3].mean(0)
    c2w = _look_at(poses[:, :3, 2].sum(0), poses[:, :3, 1].sum(0), centre)
    return np.concatenate([c2w, poses[0, :3, -1:]], 1)


def _recenter(poses):
    """Express all poses in the frame of their average (load_llff.py:162-175)."""
    out = poses + 0
    row = np.reshape([0, 0, 0, 1.0], [1, 4])
    avg = np.concatenate([_average_pose(poses)[:3, :4], row], -2)
    full = np.concatenate([poses[:, :3, :4], np.tile(row[None], [poses.shape[0], 1, 1])], -2)
    out[:, :3, :4] = (np.linalg.inv(avg) @ full)[:, :3, :4]
    return out


def _spiral_path(c2w, up, rads, focal, zrate, rots, N):
    """load_llff.py:148-158."""
    rads = np.array(list(rads) + [1.0])
    hwf = c2w[:, 4:5]
    out = []
    for theta in np.linspace(0.0, 2.0 * np.pi * rots, N + 1)[:-1]:
        c = np.dot(c2w[:3, :4], np.array([np.cos(theta), -np.sin(theta), -np.sin(theta * zrate), 1.0]) * rads)
        z = _unit(c - np.dot(c2w[:3, :4], np.array([0, 0, -focal, 1.0])))
        out.append(np.concatenate([_look_at(z, up, c), hwf], 1))
    return out


def _spherify(poses, bds):
    """360-degree captures: recentre on the point closest to all optical axes, unit mean radius, circular render
    path (load_llff.py:181-240).  `bds` is rescaled in place like the reference does."""
    def to44(p):
        return np.concatenate([p, np.tile(np.reshape(np.eye(4)[-1, :], [1, 1, 4]), [p.shape[0], 1, 1])], 1)

    d, o = poses[:, :3, 2:3], poses[:, :3, 3:4]
    A = np.eye(3) - d * np.transpose(d, [0, 2, 1])
    b = -A @ o
    centre = np.squeeze(-np.linalg.inv((np.transpose(A, [0, 2, 1]) @ A).mean(0)) @ b.mean(0))
    up = (poses[:, :3, 3] - centre).mean(0)
    v0 = _unit(up)
    v1 = _unit(np.cross([0.1, 0.2, 0.3], v0))
    v2 = _unit(np.cross(v0, v1))
    c2w = np.stack([v1, v2, v0, centre], 1)
    reset = np.linalg.inv(to44(c2w[None])) @ to44(poses[:, :3, :4])
    rad = np.sqrt(np.mean(np.sum(np.square(reset[:, :3, 3]), -1)))
    sc = 1.0 / rad
    reset[:, :3, 3] *= sc
    bds *= sc
    rad *= sc
    zh = np.mean(reset[:, :3, 3], 0)[2]
    rc = np.sqrt(rad ** 2 - zh ** 2)
    path = []
    for th in np.linspace(0.0, 2.0 * np.pi, 120):
        cam = np.array([rc * np.cos(th), rc * np.sin(th), zh])
        z = _unit(cam)
        x = _unit(np.cross(z, np.array([0, 0, -1.0])))
        y = _unit(np.cross(z, x))
        path.append(np.stack([x, y, z, cam], 1))
    path = np.stack(path, 0)
    hwf = poses[0, :3, -1:]
    path = np.concatenate([path, np.broadcast_to(hwf, path[:, :3, -1:].shape)], -1)
    reset = np.concatenate([reset[:, :3, :4], np.broadcast_to(hwf, reset[:, :3, -1:].shape)], -1)
    return reset, path, bds


def _ensure_downsampled(basedir, factor):
    """`images_<factor>/` as PNGs.  The reference shells out to ImageMagick's mogrify (load_llff.py:8-56); here the
    directory is produced with an area-filter resize when it is missing (the LLFF downloads ship it ready-made)."""
    dst = os.path.join(basedir, "images_%d" % factor)
    if os.path.exists(dst):
        return
    src = os.path.join(basedir, "images")
    names = [f for f in sorted(os.listdir(src)) if f.endswith(_IMG_EXT)]
    os.makedirs(dst)
    import cv2
    for f in names:
        img = _imread(os.path.join(src, f))
        H, W = img.shape[0] // factor, img.shape[1] // factor
        small = _area_resize(img, W, H)
        cv2.imwrite(os.path.join(dst, os.path.splitext(f)[0] + ".png"), small[..., ::-1] if small.ndim == 3 else small)


def _load_llff_raw(basedir, factor=None, load_imgs=True):
    """poses [3,5,N] (last column = H, W, focal at the loaded resolution), bds [2,N], imgs [H,W,3,N] in [0,1]
    (load_llff.py:62-114)."""
    arr = np.load(os.path.join(basedir, "poses_bounds.npy"))
    poses = arr[:, :-2].reshape([-1, 3, 5]).transpose([1, 2, 0])
    bds = arr[:, -2:].transpose([1, 0])
    sfx = ""
    if factor is not None:
        sfx = "_%d" % factor
        _ensure_downsampled(basedir, factor)
    else:
        factor = 1
    imgdir = os.path.join(basedir, "images" + sfx)
    if not os.path.exists(imgdir):
        raise FileNotFoundError("%s does not exist" % imgdir)
    files = [os.path.join(imgdir, f) for f in sorted(os.listdir(imgdir)) if f.endswith(("JPG", "jpg", "png"))]
    if poses.shape[-1] != len(files):
        raise ValueError("mismatch between %d images and %d poses in %s" % (len(files), poses.shape[-1], basedir))
    sh = _imread(files[0]).shape
    poses[:2, 4, :] = np.array(sh[:2]).reshape([2, 1])
    poses[2, 4, :] = poses[2, 4, :] * 1.0 / factor
    if not load_imgs:
        return poses, bds
    imgs = np.stack([_imread(f)[..., :3] / 255.0 for f in files], -1)
    return poses, bds, imgs


def load_llff_data(basedir, factor=8, recenter=True, bd_factor=0.75, spherify=False, path_zflat=False):
    """-> images [N,H,W,3], poses [N,3,5], bds [N,2], render_poses [M,3,5], i_test (load_llff.py:243-319)."""
    poses, bds, imgs = _load_llff_raw(basedir, factor=factor)
    # LLFF stores [down, right, back]; NeRF wants [right, up, back]; image index first
    poses = np.concatenate([poses[:, 1:2, :], -poses[:, 0:1, :], poses[:, 2:, :]], 1)
    poses = np.moveaxis(poses, -1, 0).astype(np.float32)
    images = np.moveaxis(imgs, -1, 0).astype(np.float32)
    bds = np.moveaxis(bds, -1, 0).astype(np.float32)
    sc = 1.0 if bd_factor is None else 1.0 / (bds.min() * bd_factor)     # nearest depth -> 1 / bd_factor
    poses[:, :3, 3] *= sc
    bds *= sc
    if recenter:
        poses = _recenter(poses)
    if spherify:
        poses, render_poses, bds = _spherify(poses, bds)
    else:
        c2w = _average_pose(poses)
        up = _unit(poses[:, :3, 1].sum(0))
        close, inf = bds.min() * 0.9, bds.max() * 5.0
        dt = 0.75
        focal = 1.0 / ((1.0 - dt) / close + dt / inf)                     # "focus depth" of the spiral
        rads = np.percentile(np.abs(poses[:, :3, 3]), 90, 0)
        n_views, n_rots = 120, 2
        if path_zflat:
            c2w[:3, 3] = c2w[:3, 3] + (-close * 0.1) * c2w[:3, 2]
            rads[2] = 0.0
            n_rots, n_views = 1, n_views // 2
        render_poses = _spiral_path(c2w, up, rads, focal, zrate=0.5, rots=n_rots, N=n_views)
    render_poses = np.array(render_poses).astype(np.float32)
    c2w = _average_pose(poses)
    i_test = np.argmin(np.sum(np.square(c2w[:3, 3] - poses[:, :3, 3]), -1))     # view closest to the average pose
    return images.astype(np.float32), poses.astype(np.float32), bds, render_poses, i_test


# ------------------------------------------------------------------------------------------ DeepVoxels
def load_dv_data(scene="cube", basedir="/data/deepvoxels", testskip=8):
    """-> imgs, poses [N,3,4], render_poses, [512,512,focal], i_split (load_deepvoxels.py:6-110)."""
    H = W = 512
    root = "{}/train/{}/".format(basedir, scene)
    with open(os.path.join(root, "intrinsics.txt")) as f:
        f0 = float(f.readline().split()[0])
        f.readline()                       # grid barycenter
        f.readline()                       # near plane
        f.readline()                       # scale
        height, _width = map(float, f.readline().split())
    focal = H / height * f0

    def poses_in(d):
        files = [os.path.join(d, n) for n in sorted(os.listdir(d)) if n.endswith("txt")]
        p = np.stack([np.array([float(x) for x in open(n).read().split()]).reshape([4, 4]).astype(np.float32)
                      for n in files], 0)
        flip = np.array([[1, 0, 0, 0], [0, -1, 0, 0], [0, 0, -1, 0], [0, 0, 0, 1.0]])
        return (p @ flip)[:, :3, :4].astype(np.float32)

    def images_in(d, skip):
        files = [n for n in sorted(os.listdir(d)) if n.endswith("png")]
        return np.stack([_imread(os.path.join(d, n)) / 255.0 for n in files[::skip]], 0).astype(np.float32)

    poses = poses_in(os.path.join(root, "pose"))
    test_p = poses_in("{}/test/{}/pose".format(basedir, scene))[::testskip]
    val_p = poses_in("{}/validation/{}/pose".format(basedir, scene))[::testskip]
    parts = [images_in(os.path.join(root, "rgb"), 1), images_in("{}/validation/{}/rgb".format(basedir, scene), testskip),
             images_in("{}/test/{}/rgb".format(basedir, scene), testskip)]
    counts = np.cumsum([0] + [x.shape[0] for x in parts])
    i_split = [np.arange(counts[i], counts[i + 1]) for i in range(3)]
    return np.concatenate(parts, 0), np.concatenate([poses, val_p, test_p], 0), test_p, [H, W, focal], i_split


# ------------------------------------------------------------------------------------------ synthetic scenes
def synthetic_blender(H=400, W=400, n_images=8, seed=0):
    """Blender-shaped synthetic scene: camera_angle_x = 0.6911, poses on the
    pose_spherical(theta, -30, 4) orbit, uniform-random target images."""
    rng = np.random.default_rng(seed)
    focal = 0.5 * W / np.tan(0.5 * 0.6911112070083618)
    K = np.array([[focal, 0, 0.5 * W], [0, focal, 0.5 * H], [0, 0, 1]])
    poses = np.stack([_pose_spherical(float(t), -30.0, 4.0) for t in rng.uniform(-180, 180, n_images)], 0)
    images = np.random.default_rng(seed + 1).random((n_images, H, W, 3), dtype=np.float32)
    return images, poses.astype(np.float32), [H, W, focal], K


def synthetic_llff(H=378, W=504, n_images=4, seed=0):
    """LLFF-shaped synthetic scene (configs/fern.txt at factor 8): focal = 3260.5 / 8,
    forward-facing poses = identity rotation perturbed by small rotations, N(0, 0.05^2) translations."""
    rng = np.random.default_rng(seed)
    focal = 3260.5 / 8.0 * (W / 504.0)
    K = np.array([[focal, 0, 0.5 * W], [0, focal, 0.5 * H], [0, 0, 1]])
    poses = []
    for _ in range(n_images):
        w = rng.normal(0.0, 0.02, 3)                       # small rotation vector
        th = np.linalg.norm(w) + 1e-12
        k = w / th
        Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
        R = np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * (Kx @ Kx)
        c2w = np.eye(4)
        c2w[:3, :3] = R
        c2w[:3, 3] = rng.normal(0.0, 0.05, 3)
        poses.append(c2w)
    images = np.random.default_rng(seed + 1).random((n_images, H, W, 3), dtype=np.float32)
    return images, np.stack(poses, 0).astype(np.float32), [H, W, focal], K


# ------------------------------------------------------------------------------------------ dispatch
def load_datasets(args):
    """-> images, poses, render_poses, hwf, i_split, K, bds_dict (utils.py:216-313)."""
    kind = getattr(args, "dataset_type", "synthetic")
    K = None
    if kind == "llff":
        images, poses, bds, render_poses, i_test = load_llff_data(args.datadir, args.factor, recenter=True,
                                                                   bd_factor=0.75, spherify=args.spherify)
        hwf = poses[0, :3, -1]
        poses = poses[:, :3, :4]
        print("Loaded llff", images.shape, render_poses.shape, hwf, args.datadir)
        i_test = [i_test] if not isinstance(i_test, list) else i_test
        if args.llffhold > 0:
            print("Auto LLFF holdout,", args.llffhold)
            i_test = np.arange(images.shape[0])[::args.llffhold]
        i_val = i_test
        i_train = np.array([i for i in np.arange(int(images.shape[0])) if (i not in i_test and i not in i_val)])
        if args.no_ndc:
            near, far = np.ndarray.min(bds) * 0.9, np.ndarray.max(bds) * 1.0
        else:
            near, far = 0.0, 1.0
        print("NEAR FAR", near, far)
    elif kind in ("blender", "LINEMOD"):
        if kind == "blender":
            images, poses, render_poses, hwf, i_split, near, far = load_blender_data(args.datadir, args.half_res,
                                                                                      args.testskip)
        else:
            images, poses, render_poses, hwf, K, i_split, near, far = load_LINEMOD_data(args.datadir, args.half_res,
                                                                                         args.testskip)
        print("Loaded", kind, images.shape, render_poses.shape, hwf, args.datadir)
        i_train, i_val, i_test = i_split
        if args.white_bkgd:        # composite RGBA over white (utils.py:258-261)
            images = images[..., :3] * images[..., -1:] + (1.0 - images[..., -1:])
        else:
            images = images[..., :3]
    elif kind == "deepvoxels":
        images, poses, render_poses, hwf, i_split = load_dv_data(scene=args.shape, basedir=args.datadir,
                                                                 testskip=args.testskip)
        print("Loaded deepvoxels", images.shape, render_poses.shape, hwf, args.datadir)
        i_train, i_val, i_test = i_split
        hemi_R = np.mean(np.linalg.norm(poses[:, :3, -1], axis=-1))
        near, far = hemi_R - 1.0, hemi_R + 1.0
    elif kind == "synthetic":
        H = W = 400 if not getattr(args, "half_res", False) else 200
        images, poses, hwf, K = synthetic_blender(H, W)
        n = images.shape[0]
        i_train, i_val, i_test = np.arange(0, n - 2), np.arange(n - 2, n - 1), np.arange(n - 1, n)
        render_poses = np.stack([_pose_spherical(a, -30.0, 4.0) for a in np.linspace(-180, 180, 8 + 1)[:-1]], 0)
        near, far = 2.0, 6.0
    else:
        print("Unknown dataset type", kind, "exiting")
        return None

    H, W, focal = hwf
    H, W = int(H), int(W)
    hwf = [H, W, focal]
    if K is None:
        K = np.array([[focal, 0, 0.5 * W], [0, focal, 0.5 * H], [0, 0, 1]])
    if getattr(args, "render_test", False):
        render_poses = np.array(poses[i_test])
    return images, poses, render_poses, hwf, (i_train, i_val, i_test), K, {"near": near, "far": far}
