"""Calibration data the observation path consumes (values verbatim from the reference).

These are measured numbers, not code: sim->GS similarity transforms, label maps, per-actor fine-tuning
offsets/scales, scan/task joint positions and camera calibration.  Source of every table is cited;
only the subset the hot path and BASELINE configs 1-4 touch is carried.

Reference: gsworld/constants.py (sim2gs_*:30-49,181-342; object_offset/scale:345-398;
*_gs_semantics:402-468; qpos:51-103,163-179; camera calib:507-539).
"""
from __future__ import annotations

import numpy as np

_f = np.float32


def _m(rows):
    return np.array(rows, dtype=_f)


# x-axis half turn used by a few object calibrations (constants.py:14-18)
_X180 = np.array([[1, 0, 0], [0, -1, 0], [0, 0, -1]])

# ---- robot sim->GS similarity [s*R | t] (constants.py:30-49) ----
SIM2GS_ROBOT = {
    "fr3": _m([
        [0.65203872, 0.70075277, 0.03073432, -0.08619287],
        [0.03194594, 0.01225097, -0.95706996, -0.75944751],
        [-0.70069858, 0.65264769, -0.01503433, 0.25320947],
        [0.0, 0.0, 0.0, 1.0]]),
    "xarm": _m([
        [-0.97002696, 0.2247966, 0.10835464, 0.32787871],
        [0.05080531, 0.60369423, -0.7976206, 0.37823396],
        [-0.24432164, -0.76697216, -0.59605971, 0.45637834],
        [0.0, 0.0, 0.0, 1.0]]),
    "r1": _m([
        [2.70573310e-01, -8.82001081e-01, 3.56843272e-03, -1.01723192e-02],
        [-2.38923961e-04, -3.80586011e-03, -9.22569247e-01, -6.73930139e-02],
        [8.82008267e-01, 2.70570074e-01, -1.34459800e-03, -1.45273889e-02],
        [0.0, 0.0, 0.0, 1.0]]),
}


def robot_family(scene_gs_cfg_name: str) -> str:
    """Same substring dispatch as GSWorldWrapper.__init__ (gs_world_wrapper.py:52-62)."""
    if "xarm" in scene_gs_cfg_name:
        return "xarm"
    if "fr3" in scene_gs_cfg_name or "franka" in scene_gs_cfg_name:
        return "fr3"
    if "r1" in scene_gs_cfg_name:
        return "r1"
    raise NotImplementedError(scene_gs_cfg_name)


# ---- object sim->GS similarity (constants.py:181-342) ----
_bread = _m([
    [3.98912, 10.5397, 0.649514, 0.345585],
    [-7.51925, 2.34766, 8.08519, 0.941237],
    [7.41409, -3.28989, 7.8504, 1.70994],
    [0.0, 0.0, 0.0, 1.0]])
_bread[:3, :3] = _bread[:3, :3] @ _X180

_mustard = _m([
    [0.510841, -0.618467, -0.0283021, 0.000536782],
    [-0.0116977, 0.0270441, -0.802118, -0.299513],
    [0.619003, 0.51091, 0.00819852, -0.0257972],
    [0, 0, 0, 1]])
_snack = _m([
    [0.167587, -0.728182, -0.664579, 0.0772496],
    [-0.961955, -0.268324, 0.051428, -0.137463],
    [-0.215772, 0.630676, -0.745446, 0.262004],
    [0.0, 0.0, 0.0, 1.0]])
_soup = _m([
    [-0.000393405, -0.892165, -0.000564396, -0.0584909],
    [0.00982369, 0.000560062, -0.892111, -0.294318],
    [0.892111, -0.0003996, 0.00982345, -0.045191],
    [0.0, 0.0, 0.0, 1.0]])
_green_can = _m([
    [-7.33941866e-01, 3.36139655e-02, -5.37589582e-01, 4.31403401e-02],
    [-2.65421561e-02, -9.09764955e-01, -2.06484809e-02, -2.37379279e-01],
    [-5.37985103e-01, -9.73211127e-04, 7.34420997e-01, -1.01842246e-01],
    [0.0, 0.0, 0.0, 1.0]])

SIM2GS_OBJECT = {
    "005_tomato_soup_can": _soup,
    "006_mustard_bottle": _mustard,
    "snack_box": _snack,
    "bread_slice": _bread,
    "dtc_green_can": _green_can,
}

# ---- per-actor fine-tuning (constants.py:345-398) ----
OBJECT_OFFSET = {
    "005_tomato_soup_can": [0.0, 0.0, 0.01],
    "006_mustard_bottle": [0.0, 0.0, 0.02],
    "snack_box": [0.0, 0.0, 0.035],
    "bread_slice": [0.0, 0.0, 0.04],
    "dtc_green_can": [-0.04, 0.0, -0.03],
    "xarm_arm": [0.0, 0.0, 0.05],
}
OBJECT_SCALE = {
    "005_tomato_soup_can": 1,
    "006_mustard_bottle": 1,
    "snack_box": 1,
    "bread_slice": 0.95,
    "dtc_green_can": 1,
}

# ---- labels; -1 is background (constants.py:401-468) ----
LINK_LABELS = {
    "fr3": {
        "base": 0, "fr3_link0": 1, "fr3_link1": 2, "fr3_link2": 3, "fr3_link3": 4, "fr3_link4": 5,
        "fr3_link5": 6, "fr3_link6": 7, "fr3_link7": 8, "fr3_link8": 9, "fr3_hand": [10, 14, 15],
        "fr3_hand_tcp": 11, "fr3_leftfinger": 12, "fr3_rightfinger": 13,
    },
    "xarm": {
        "world": 0, "link_base": 1, "link1": 2, "link2": 3, "link3": 4, "link4": 5, "link5": 6,
        "link6": [7, 8], "xarm_gripper_base_link": 9, "left_outer_knuckle": 10, "left_inner_knuckle": 11,
        "right_outer_knuckle": 12, "right_inner_knuckle": 13, "xarm_hand_tcp": 14, "left_finger": 15,
        "right_finger": 16,
    },
}
OBJECT_LABELS = {
    "006_mustard_bottle": 100, "snack_box": 103, "005_tomato_soup_can": 110, "bread_slice": 111,
    "dtc_green_can": 201,
}

# ---- joint positions (constants.py:51-103,163-179) ----
_fr3_scan = np.array([0.00879998, -0.62698067, -0.00591884, -2.25830053, -0.00557862, 1.63165594,
                      0.78539816, 4.04319502e-02, 4.04319502e-02, 0.02], dtype=_f)
_fr3_task = _fr3_scan[:9].copy()
_xarm_task = np.array([0.0, 0.0, -np.pi / 3, 0.0, np.pi / 3, 0.0, 0, 0, 0, 0, 0, 0], dtype=_f)
ROBOT_SCAN_QPOS = {"fr3_umi": _fr3_scan, "xarm6_uf_gripper": _xarm_task}
ROBOT_TASK_INIT_QPOS = {"fr3_umi": _fr3_task, "xarm6_uf_gripper": _xarm_task}

# ---- cameras (constants.py:507-539): OpenCV-convention camera->parent matrices + RealSense K ----
RS_D435I_RGB_K = _m([[606.12145996, 0.0, 318.3548584], [0.0, 605.1428833, 242.92498779], [0.0, 0.0, 1.0]])
RS_D435I_RES = (640, 480)  # (W, H) the K above is calibrated for (real_fr3_env.py:47-52)
WRIST2EEF = _m([
    [0.00561756, -0.99991452, 0.01180684, 0.0691971],
    [0.99993738, 0.00573118, 0.00961197, 0.02580245],
    [-0.00967881, 0.01175211, 0.9998841, -0.1056441],
    [0.0, 0.0, 0.0, 1.0]])
RIGHT2BASE = _m([
    [-0.025185470710454363, 0.9003537485256276, -0.43442930331751733, 0.8003658631290567],
    [0.9990845637502204, 0.007637667199582072, -0.04209157297821219, 0.014761293894194942],
    [-0.034579279071787865, -0.4350917070636938, -0.8997218903101533, 0.8497237283025128],
    [0.0, 0.0, 0.0, 1.0]])
XARM_RIGHT2BASE = _m([
    [-0.99815940, 0.02312000, 0.05609515, 0.38209513],
    [-0.00610404, 0.88159275, -0.47197380, 0.40018010],
    [-0.06036488, -0.47144645, -0.87982790, 0.46095666],
    [0.0, 0.0, 0.0, 1.0]])
XARM_WRIST2BASE = _m([
    [-0.0375638, -0.9982628, -0.04539683, 0.01998455],
    [0.99928665, -0.03734544, -0.00564907, -0.00621691],
    [0.00394388, -0.04557664, 0.99895304, -0.0705968],
    [0.0, 0.0, 0.0, 1.0]])
