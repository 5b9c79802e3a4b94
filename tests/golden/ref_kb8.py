"""Python transliteration of KannalaBrandt8::unproject / project / TriangulateMatches / Triangulate
(src/geometry/cameramodels/kannalabrandt8.cpp:11-27, 82-109, 324-393, 485-498) with the REAL cv::SVD::compute of this
container's OpenCV (cv2.SVDecomp, 4.13) for the 4x4 null vector, used to generate tests/golden/kb8.npz.  Scalars are
np.float32 wherever the reference computes in float, Python floats (doubles) where it computes in double."""
import math

import cv2
import numpy as np

f32 = np.float32
PI_2 = f32(math.pi / 2.0)


def unproject(p, prm, precision=1e-6):
    pwx = f32((f32(p[0]) - prm[2]) / prm[0])
    pwy = f32((f32(p[1]) - prm[3]) / prm[1])
    scale = f32(1.0)
    theta_d = f32(np.sqrt(f32(pwx * pwx + pwy * pwy)))
    theta_d = min(max(-PI_2, theta_d), PI_2)
    if theta_d > 1e-8:
        theta = theta_d
        for _ in range(10):
            theta2 = f32(theta * theta)
            theta4 = f32(theta2 * theta2)
            theta6 = f32(theta4 * theta2)
            theta8 = f32(theta4 * theta4)
            k0t2, k1t4, k2t6, k3t8 = f32(prm[4] * theta2), f32(prm[5] * theta4), f32(prm[6] * theta6), f32(prm[7] * theta8)
            num = f32(f32(theta * f32(f32(f32(f32(f32(1) + k0t2) + k1t4) + k2t6) + k3t8)) - theta_d)
            den = f32(f32(f32(f32(f32(1) + f32(f32(3) * k0t2)) + f32(f32(5) * k1t4)) + f32(f32(7) * k2t6)) + f32(f32(9) * k3t8))
            theta_fix = f32(num / den)
            theta = f32(theta - theta_fix)
            if abs(theta_fix) < precision:
                break
        scale = f32(f32(np.tan(theta)) / theta_d)
    return np.array([pwx * scale, pwy * scale, f32(1.0)], f32)


def project(p3, prm):
    x, y, z = f32(p3[0]), f32(p3[1]), f32(p3[2])
    x2y2 = f32(f32(x * x) + f32(y * y))
    theta = f32(np.arctan2(f32(np.sqrt(x2y2)), z))
    psi = f32(np.arctan2(y, x))
    t2 = f32(theta * theta)
    t3 = f32(theta * t2)
    t5 = f32(t3 * t2)
    t7 = f32(t5 * t2)
    t9 = f32(t7 * t2)
    r = f32(f32(f32(f32(theta + f32(prm[4] * t3)) + f32(prm[5] * t5)) + f32(prm[6] * t7)) + f32(prm[7] * t9))
    return (f32(f32(f32(prm[0] * r) * f32(np.cos(psi))) + prm[2]), f32(f32(f32(prm[1] * r) * f32(np.sin(psi))) + prm[3]))


def triangulate(p1, p2, T1, T2):
    A = np.zeros((4, 4), f32)
    A[0] = f32(p1[0]) * T1[2] - T1[0]
    A[1] = f32(p1[1]) * T1[2] - T1[1]
    A[2] = f32(p2[0]) * T2[2] - T2[0]
    A[3] = f32(p2[1]) * T2[2] - T2[1]
    _, _, vt = cv2.SVDecomp(A, flags=cv2.SVD_MODIFY_A | cv2.SVD_FULL_UV)
    x = vt[3].astype(f32)
    return (x[:3] / x[3]).astype(f32)


def triangulate_matches(prm1, prm2, kp1, kp2, R12, t12, sigma_level, unc):
    """-> (return value, p3D or None)"""
    r1, r2 = unproject(kp1, prm1), unproject(kp2, prm2)
    r21 = (R12.astype(f32) @ r2).astype(f32)
    cos_par = f32(float(np.dot(r1.astype(np.float64), r21.astype(np.float64))) /
                  (float(np.linalg.norm(r1.astype(np.float64))) * float(np.linalg.norm(r21.astype(np.float64)))))
    if cos_par > 0.9998:
        return f32(-1), None
    T1 = np.hstack([np.eye(3, dtype=f32), np.zeros((3, 1), f32)])
    R21 = R12.T.astype(f32)
    t21 = (-(R21 @ t12.astype(f32))).astype(f32)
    T2 = np.hstack([R21, t21.reshape(3, 1)]).astype(f32)
    x3D = triangulate((r1[0], r1[1]), (r2[0], r2[1]), T1, T2)
    z1 = x3D[2]
    if z1 <= 0:
        return f32(-1), None
    z2 = f32(float(np.dot(R21[2].astype(np.float64), x3D.astype(np.float64))) + float(t21[2]))
    if z2 <= 0:
        return f32(-1), None
    u1 = project(x3D, prm1)
    ex, ey = f32(u1[0] - f32(kp1[0])), f32(u1[1] - f32(kp1[1]))
    if float(f32(f32(ex * ex) + f32(ey * ey))) > 5.991 * float(sigma_level):
        return f32(-1), None
    x3D2 = ((R21 @ x3D).astype(f32) + t21).astype(f32)
    u2 = project(x3D2, prm2)
    ex, ey = f32(u2[0] - f32(kp2[0])), f32(u2[1] - f32(kp2[1]))
    if float(f32(f32(ex * ex) + f32(ey * ey))) > 5.991 * float(unc):
        return f32(-1), None
    return z1, x3D
