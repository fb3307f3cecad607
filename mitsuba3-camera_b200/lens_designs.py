"""Lens prescriptions for the `realisticlens` sensor: the data side of
src/sensors/realisticlens.cpp:844-897 (design selection) and :936-1566 (the 14 hard-coded tables).

The reference hard-codes its prescriptions in C++; here they are DATA (data/lens_presets.json,
extracted by tools/extract_presets.py) plus three parametric constructions (singlet, doublet,
hypercentric) restated below in float32. Everything is turned into flat float32/int32 arrays that
the C ABI (include/lensrt.h: lrt_lens_from_prescription / lrt_lens_from_interfaces) consumes; the
float32 conversion of the tables to metres and the thick-lens autofocus happen in C++ behind it.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field

import numpy as np

f32 = np.float32
KIND = {"sphere": 0, "plane": 1, "aperture": 2, "asphere": 3}
KIND_NAMES = {v: k for k, v in KIND.items()}
_PRESETS = None


def presets() -> dict:
    global _PRESETS
    if _PRESETS is None:
        with open(os.path.join(os.path.dirname(__file__), "data", "lens_presets.json")) as f:
            _PRESETS = json.load(f)
    return _PRESETS


def _lit(tok: str) -> np.float32:
    """One source literal -> float32 (the tables hold the decimal strings of the reference)."""
    if "/" in tok:
        a, b = tok.split("/")
        return f32(f32(float(a)) / f32(float(b)))
    return f32(float(tok))


def material_row(m: dict) -> np.ndarray:
    """[model, p0..p5]: Cauchy (0, A, B) or Sellmeier (1, B1..B3, C1..C3); realisticlens.cpp:151-167."""
    row = np.zeros(7, f32)
    if m["model"] == "cauchy":
        row[1], row[2] = _lit(m["A"]), _lit(m["B"])
    else:
        row[0] = 1
        row[1:4] = [_lit(x) for x in m["B"]]
        row[4:7] = [_lit(x) for x in m["C"]]
    return row


def material_ior(row: np.ndarray, wavelength_um=0.5893) -> np.float32:
    """realisticlens.cpp:226-240 in float32 (used only for the parametric designs' interface rows)."""
    w = f32(wavelength_um)
    w2 = f32(w * w)
    if row[0] == 0:
        return f32(row[1] + f32(row[2] / w2))
    ior = f32(1)
    for k in range(3):
        ior = f32(ior + f32(f32(row[1 + k] * w2) / f32(w2 - row[4 + k])))
    return f32(np.sqrt(ior))


@dataclass
class Prescription:
    """Front->rear table in millimetres (what build_lens_from_data takes, :1568-1575)."""
    kinds: np.ndarray        # int32 [n]
    thickness: np.ndarray    # f32 [n]
    radius: np.ndarray       # f32 [n]
    semi_diameter: np.ndarray  # f32 [n]
    kappa: np.ndarray        # f32 [n]
    ai: np.ndarray           # f32 [n,7]
    materials: np.ndarray    # f32 [n+1,7], world side first
    scale: np.float32 = f32(1)
    stop_ratio: np.float32 = f32(1)
    fine_focus: bool = True
    material_names: list = field(default_factory=list)


@dataclass
class InterfaceList:
    """Ready-made interfaces in metres, rear->front, rows of 16 floats (see include/lensrt.h)."""
    rows: np.ndarray         # f32 [n,16]
    film_z: np.float32 = f32(0)
    fine_focus: bool = True
    stop_ratio: np.float32 = f32(1)
    materials: np.ndarray | None = None      # [n,2,7] Cauchy / Sellmeier rows of every interface's (left, right) media: dispersion mode


def prescription_from_preset(name: str, stop_ratio=1.0) -> Prescription:
    d = presets()[name]
    s = d["surfaces"]
    n = len(s)
    p = Prescription(
        kinds=np.array([KIND[x["kind"]] for x in s], np.int32),
        thickness=np.array([_lit(x["thickness_mm"]) for x in s], f32),
        radius=np.array([_lit(x["radius_mm"]) for x in s], f32),
        semi_diameter=np.array([_lit(x["semi_diameter_mm"]) for x in s], f32),
        kappa=np.array([_lit(x["kappa"]) for x in s], f32),
        ai=np.array([[_lit(a) for a in x["ai"]] for x in s], f32).reshape(n, 7),
        materials=np.stack([material_row(m) for m in d["materials"]]),
        scale=_lit(d["scale"]),
        # exp2* force m_stopdown_ratio = 0.5 (:1563); all others use the `stop_ratio` property (:850)
        stop_ratio=_lit(d["stop_ratio"]) if d["stop_ratio"] else f32(stop_ratio),
        fine_focus=bool(d["fine_focus"]),
        material_names=[m["name"] for m in d["materials"]],
    )
    return p


_AIR = np.array([0, 1.000277, 0, 0, 0, 0, 0], f32)


def _row(kind, z, elem_r, curv=0.0, il=1.0, ir=1.0):
    r = np.zeros(16, f32)
    r[0], r[1], r[2], r[3], r[13], r[14] = kind, z, elem_r, curv, il, ir
    return r


NOMINAL_RGB = f32(0.5893)                          # compute_ior(Float(0.5893f)), realisticlens.cpp:170-173
NOMINAL_SPECTRAL = f32(f32(0.001) * f32(589.3))    # compute_ior(0.001f * ray.wavelengths[0]) with Wavelength(589.3f) focus rays


def singlet(object_distance, curvature_radius, lens_radius, nominal=NOMINAL_RGB) -> InterfaceList:
    """realisticlens.cpp:936-954 (biconvex singlet placed by the thin-lens equation)."""
    od, R, lr = f32(object_distance), f32(curvature_radius), f32(lens_radius)
    distance = max(od, f32(f32(4.001) * R))
    z = f32(f32(f32(0.5) * distance) * f32(f32(1) - np.sqrt(f32(f32(1) - f32(f32(f32(4) * R) / distance)))))
    q = f32(f32(lr / R) * f32(lr / R))
    thick = f32(f32(f32(2) * R) * f32(f32(1) - np.sqrt(f32(f32(1) - q))))
    glass = np.array([0, 1.5046, f32(f32(5) * f32(0.00420)), 0, 0, 0, 0], f32)
    na, ng = material_ior(_AIR, nominal), material_ior(glass, nominal)
    rows = np.stack([_row(0, z, lr, R, na, ng), _row(0, f32(z + thick), lr, -R, ng, na)])
    return InterfaceList(rows, materials=np.array([[_AIR, glass], [glass, _AIR]], f32))


def doublet(object_distance, R, lens_radius, nominal=NOMINAL_RGB) -> InterfaceList:
    """realisticlens.cpp:957-975 (plano / sphere / sphere / stop)."""
    od, R, lr = f32(object_distance), f32(R), f32(lens_radius)
    fl = f32(f32(2) * R)
    distance = max(od, f32(f32(4.001) * fl))
    z = f32(f32(f32(0.5) * distance) * f32(f32(1) - np.sqrt(f32(f32(1) - f32(f32(f32(4) * fl) / distance)))))
    q = f32(f32(lr / R) * f32(lr / R))
    thick = f32(f32(f32(2) * R) * f32(f32(1) - np.sqrt(f32(f32(1) - q))))
    g1 = np.array([0, 1.5046, 0.00420, 0, 0, 0, 0], f32)
    g2 = np.array([0, 1.5046, 0.00860948454, 0, 0, 0, 0], f32)
    na, n1, n2 = material_ior(_AIR, nominal), material_ior(g1, nominal), material_ior(g2, nominal)
    rows = np.stack([
        _row(1, z, lr, 0, na, n2),
        _row(0, f32(z + thick), lr, R, n2, n1),
        _row(0, f32(z + f32(f32(2) * thick)), lr, -R, n1, na),
        _row(2, f32(z + f32(f32(3) * thick)), f32(f32(1) * lr), 0, na, na),
    ])
    return InterfaceList(rows, materials=np.array([[_AIR, g2], [g2, g1], [g1, _AIR], [_AIR, _AIR]], f32))


def hypercentric(ap_radius=1.0, focal_lengths=2.0, nominal=NOMINAL_RGB) -> InterfaceList:
    """realisticlens.cpp:1322-1370 (telecentric / hypercentric test optic; no fine focus)."""
    apr, fls = f32(ap_radius), f32(focal_lengths)
    sf6 = np.array([0, 1.757435971209294, 0.016580283248837778, 0, 0, 0, 0], f32)
    na, ng = material_ior(_AIR, nominal), material_ior(sf6, nominal)
    target, lens_radius, focal, track, curv, lens_t = map(f32, (5.0, 120.0, 210.0, 450.0, 310.505660861, 50.0))
    z_ap_to_lens = f32(focal * fls)
    z_film_to_ap = f32(f32(z_ap_to_lens * f32(target - apr)) * f32(f32(1) / f32(lens_radius + apr)))
    curv_radii = [curv, f32(-curv), f32(np.inf)]
    thick = [lens_t, f32(z_ap_to_lens - f32(lens_t * f32(0.5))), z_film_to_ap]
    elem = [lens_radius, lens_radius, apr]
    iors = [na, ng, na, na]
    media = [_AIR, sf6, _AIR, _AIR]
    rows, z_pos, mats = [], f32(0), []
    for i in (2, 1, 0):
        z_pos = f32(z_pos + thick[i])
        if i == 2:
            rows.append(_row(2, f32(f32(0.001) * z_pos), f32(f32(0.001) * elem[i]), 0, na, na))
            mats.append([_AIR, _AIR])
        else:
            mats.append([media[i + 1], media[i]])
            rows.append(_row(0, f32(f32(0.001) * z_pos), f32(f32(0.001) * elem[i]),
                             f32(f32(-0.001) * curv_radii[i]), iors[i + 1], iors[i]))
    film_z = f32(f32(0.001) * f32(track - f32(f32(thick[0] + thick[1]) + thick[2])))
    return InterfaceList(np.stack(rows), film_z=film_z, fine_focus=False, materials=np.array(mats, f32))


PARAMETRIC = ("singlet", "doublet", "hypercentric")


def design(lens_design: str, props: dict | None = None):
    """realisticlens.cpp:844-897: `lens_design` string + properties -> Prescription | InterfaceList."""
    props = props or {}
    od = props.get("object_distance", 6.0)
    fl = props.get("lens_focal_length", 0.05)
    ld = props.get("lens_diameter", 0.01)
    sr = props.get("stop_ratio", 1.0)
    nominal = NOMINAL_SPECTRAL if props.get("spectral", False) else NOMINAL_RGB
    if lens_design == "doublet":
        return doublet(od, f32(f32(fl) / f32(2)), f32(f32(ld) / f32(2)), nominal)
    if lens_design == "hypercentric":
        return hypercentric(props.get("ap_size", 1.0), props.get("ap_offset", 2.0), nominal)
    if lens_design in presets():
        return prescription_from_preset(lens_design, sr)
    # "singlet" and every unknown name (:895-897)
    return singlet(od, f32(fl), f32(f32(ld) / f32(2)), nominal)


def all_designs():
    return ["singlet", "doublet", "hypercentric"] + list(presets().keys())
