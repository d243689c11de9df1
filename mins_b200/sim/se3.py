"""SO(3)/SE(3) helpers for the closed-loop harness (numpy, fp64). Conventions follow ov_core's quat_ops.h
(thirdparty/open_vins/ov_core/src/utils/quat_ops.h): JPL quaternions, R = R_GtoI, error state R_true = exp_so3(-dth) R_hat."""
import numpy as np


def skew(w):
    return np.array([[0.0, -w[2], w[1]], [w[2], 0.0, -w[0]], [-w[1], w[0], 0.0]])


def exp_so3(w):
    """quat_ops.h:231 exp_so3 (Rodrigues)."""
    w = np.asarray(w, np.float64)
    th = np.linalg.norm(w)
    K = skew(w)
    if th < 1e-7:
        return np.eye(3) + K + 0.5 * K @ K
    return np.eye(3) + np.sin(th) / th * K + (1.0 - np.cos(th)) / (th * th) * K @ K


def log_so3(R):
    """quat_ops.h log_so3: rotation vector of R."""
    c = 0.5 * (np.trace(R) - 1.0)
    c = min(1.0, max(-1.0, c))
    th = np.arccos(c)
    v = np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    if th < 1e-7:
        return 0.5 * v
    if np.pi - th < 1e-6:  # near pi: use the symmetric part
        A = 0.5 * (R + np.eye(3))
        ax = np.sqrt(np.maximum(np.diag(A), 0.0))
        k = int(np.argmax(ax))
        ax = A[:, k] / ax[k]
        if ax @ v < 0:
            ax = -ax
        return th * ax / np.linalg.norm(ax)
    return th / (2.0 * np.sin(th)) * v


def Jl_so3(w):
    """left Jacobian of SO(3) (quat_ops.h Jl_so3)."""
    w = np.asarray(w, np.float64)
    th = np.linalg.norm(w)
    K = skew(w)
    if th < 1e-7:
        return np.eye(3) + 0.5 * K + K @ K / 6.0
    return np.eye(3) + (1.0 - np.cos(th)) / th**2 * K + (th - np.sin(th)) / th**3 * K @ K


def Jr_so3(w):
    return Jl_so3(-np.asarray(w, np.float64))


def quat_jpl_to_rot(q):
    """quat_ops.h quat_2_Rot: JPL quaternion [x, y, z, w] -> rotation matrix (R_GtoI for q_GtoI)."""
    q = np.asarray(q, np.float64)
    q = q / np.linalg.norm(q)
    v, w = q[:3], q[3]
    return (2 * w * w - 1) * np.eye(3) - 2 * w * skew(v) + 2 * np.outer(v, v)


def rot_to_quat_jpl(R):
    """quat_ops.h rot_2_quat."""
    T = np.trace(R)
    q = np.zeros(4)
    if R[0, 0] >= T and R[0, 0] >= R[1, 1] and R[0, 0] >= R[2, 2]:
        q[0] = np.sqrt((1 + 2 * R[0, 0] - T) / 4)
        q[1] = (R[0, 1] + R[1, 0]) / (4 * q[0])
        q[2] = (R[0, 2] + R[2, 0]) / (4 * q[0])
        q[3] = (R[1, 2] - R[2, 1]) / (4 * q[0])
    elif R[1, 1] >= T and R[1, 1] >= R[0, 0] and R[1, 1] >= R[2, 2]:
        q[1] = np.sqrt((1 + 2 * R[1, 1] - T) / 4)
        q[0] = (R[0, 1] + R[1, 0]) / (4 * q[1])
        q[2] = (R[1, 2] + R[2, 1]) / (4 * q[1])
        q[3] = (R[2, 0] - R[0, 2]) / (4 * q[1])
    elif R[2, 2] >= T and R[2, 2] >= R[0, 0] and R[2, 2] >= R[1, 1]:
        q[2] = np.sqrt((1 + 2 * R[2, 2] - T) / 4)
        q[0] = (R[0, 2] + R[2, 0]) / (4 * q[2])
        q[1] = (R[1, 2] + R[2, 1]) / (4 * q[2])
        q[3] = (R[0, 1] - R[1, 0]) / (4 * q[2])
    else:
        q[3] = np.sqrt((1 + T) / 4)
        q[0] = (R[1, 2] - R[2, 1]) / (4 * q[3])
        q[1] = (R[2, 0] - R[0, 2]) / (4 * q[3])
        q[2] = (R[0, 1] - R[1, 0]) / (4 * q[3])
    if q[3] < 0:
        q = -q
    return q / np.linalg.norm(q)


def exp_se3(xi):
    """quat_ops.h exp_se3: xi = [w, u] -> 4x4."""
    w, u = np.asarray(xi[:3], np.float64), np.asarray(xi[3:], np.float64)
    T = np.eye(4)
    T[:3, :3] = exp_so3(w)
    T[:3, 3] = Jl_so3(w) @ u
    return T


def log_se3(T):
    w = log_so3(T[:3, :3])
    u = np.linalg.solve(Jl_so3(w), T[:3, 3])
    return np.concatenate([w, u])


def inv_se3(T):
    Ti = np.eye(4)
    Ti[:3, :3] = T[:3, :3].T
    Ti[:3, 3] = -T[:3, :3].T @ T[:3, 3]
    return Ti
