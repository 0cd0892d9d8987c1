"""Shared inputs for the isInFrustum tests: a camera pose, and map points scattered so that every rejection branch of
Frame::isInFrustum (behind the camera, outside the image, too near / too far, oblique normal) is taken."""
import numpy as np

from ar_orbslam2_b200._abi import KEYPOINT_DTYPE

FX, FY, CX, CY, MBF = 435.2, 435.2, 367.2, 252.2, 47.9     # Examples/Stereo/EuRoC.yaml:8-11,25
W, H = 752, 480
LOGS = float(np.log(np.float32(1.2)).astype(np.float32))


def pose(seed):
    rng = np.random.default_rng(seed)
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
    th = 0.3 * rng.normal()
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K
    T = np.eye(4, dtype=np.float32)
    T[:3, :3] = R.astype(np.float32)
    T[:3, 3] = rng.normal(0, 0.5, 3).astype(np.float32)
    return T


def points(seed, n=4000):
    rng = np.random.default_rng(seed + 100)
    wp = rng.normal(0, 4, (n, 3)).astype(np.float32)
    wp[:, 2] += 3
    nr = rng.normal(size=(n, 3)); nr /= np.linalg.norm(nr, axis=1, keepdims=True)
    # half of the normals point roughly back to the origin region so that the viewing-angle test passes often
    nr[: n // 2] = wp[: n // 2] / np.linalg.norm(wp[: n // 2], axis=1, keepdims=True) + 0.3 * nr[: n // 2]
    nr /= np.linalg.norm(nr, axis=1, keepdims=True)
    d = np.linalg.norm(wp, axis=1).astype(np.float32)
    maxd = (d * rng.uniform(0.6, 3.0, n)).astype(np.float32)
    mind = (maxd / np.float32(1.2 ** 7) * rng.uniform(0.5, 1.5, n)).astype(np.float32)
    consider = (rng.random(n) > 0.1).astype(np.uint8)
    return wp, nr.astype(np.float32), maxd, mind, consider


def dummy_frame(backend):
    from ar_orbslam2_b200.tracking import Frame
    k = np.zeros(4, KEYPOINT_DTYPE)
    k["x"], k["y"] = [10, 20, 30, 40], [10, 20, 30, 40]
    sc = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    return Frame(k, np.zeros((4, 32), np.uint8), sc, (0, W, 0, H), backend=backend)


def cam_dict(cam15):
    return dict(Rcw=cam15[:9], tcw=cam15[9:12], Ow=cam15[12:15], fx=FX, fy=FY, cx=CX, cy=CY, min_x=0, max_x=W, min_y=0, max_y=H, mbf=MBF,
                log_scale_factor=LOGS)
