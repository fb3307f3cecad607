#!/usr/bin/env python3
"""Extract the lens PRESCRIPTION DATA (numbers only) of the reference's hard-coded lens designs
(src/sensors/realisticlens.cpp:977-1566) into a neutral JSON table.

Run in the build container (needs /root/reference):  python tools/extract_presets.py
Output: mitsuba3-camera_b200/data/lens_presets.json  (committed; the product and the tests read it).

Format per design: {"scale", "stop_ratio", "fine_focus", "surfaces": [front->rear rows
{"kind","radius_mm","thickness_mm","semi_diameter_mm","kappa","ai":[7]}], "materials": [n+1 rows,
world side first: {"name","model":"cauchy","A","B"} | {"model":"sellmeier","B":[3],"C":[3]}]}.
Literals are stored as the decimal strings of the source so that float32 rounding is done once, by
the consumer.
"""
import json, re, sys, os

SRC = "/root/reference/src/sensors/realisticlens.cpp"
OUT = os.path.join(os.path.dirname(__file__), "..", "mitsuba3-camera_b200", "data", "lens_presets.json")

def num(tok):
    tok = tok.strip()
    if "*" in tok:   # one table entry is a float32 product ("24.57f * 0.5f", :1125)
        import numpy as np
        a, b = (np.float32(float(num(t))) for t in tok.split("*"))
        return repr(float(np.float32(a * b)))
    if tok.endswith("f") and not tok.lower().startswith("0x"):
        tok = tok[:-1]
    float(tok)  # validate
    return tok

def body_of(text, fname):
    m = re.search(r"void\s+" + fname + r"\s*\([^)]*\)\s*\{", text)
    i = m.end(); depth = 1
    while depth:
        c = text[i]
        depth += (c == "{") - (c == "}")
        i += 1
    return text[m.end():i - 1]

def strip_comments(s):
    return re.sub(r"//[^\n]*", "", s)

def vec(body, name):
    m = re.search(name + r"\s*=\s*\{(.*?)\};", body, re.S)
    inner = m.group(1).strip()
    return [t for t in (x.strip() for x in inner.replace("\n", " ").split(",")) if t]

def parse(body, subs=None, ai_override=None):
    body = strip_comments(body)
    for k, v in (subs or {}).items():
        body = re.sub(r"\b" + k + r"\b", v, body)
    mats = {}
    # Cauchy: Name = DispersiveMaterial<...>("label", A, B)   Sellmeier: ("label", Vector3f(..), Vector3f(..))
    for m in re.finditer(r"(\w+)\s*=\s*DispersiveMaterial<Float,\s*Spectrum>\(", body):
        i = m.end(); depth = 1
        while depth:
            depth += (body[i] == "(") - (body[i] == ")")
            i += 1
        inner = " ".join(body[m.end():i - 1].split())
        var = m.group(1)
        label, args = re.match(r'"([^"]+)"\s*,\s*(.*)', inner).groups()
        if "Vector3f" in args:
            vs = re.findall(r"Vector3f\(([^)]*)\)", args)
            mats[var] = {"name": label, "model": "sellmeier", "B": [num(t) for t in vs[0].split(",")],
                         "C": [num(t) for t in vs[1].split(",")]}
        else:
            a, b = args.split(",")
            mats[var] = {"name": label, "model": "cauchy", "A": num(a), "B": num(b)}
    kinds = [t.split("::")[1] for t in vec(body, "interfaceTypes")]
    n = len(kinds)
    curv = [num(t) for t in vec(body, "curv_radii")]
    thick = [num(t) for t in vec(body, "thicknesses")]
    elem = [num(t) for t in vec(body, "elem_radii")]
    kap = vec(body, "kappas")
    kap = [num(t) for t in kap] if kap else ["0.0"] * n
    m = None if re.search(r"ai_list\s*=\s*\{\s*\};", body) else \
        re.search(r"ai_list\s*=\s*\{(.*?)\n\s*\};", body, re.S)
    ai = []
    if m and m.group(1).strip():
        for row in re.findall(r"\{([^{}]*)\}", m.group(1)):
            ai.append([num(t) for t in row.split(",") if t.strip()])
    else:
        ai = [["0.0"] * 7 for _ in range(n)]
    for idx, row in (ai_override or {}).items():
        ai[idx] = row
    m = re.search(r"mats\s*=\s*\{(.*?)\};", body, re.S)
    mat_names = [t.strip() for t in m.group(1).split(",") if t.strip()]
    assert len(curv) == len(thick) == len(elem) == len(kap) == len(ai) == n and len(mat_names) == n + 1, \
        (n, len(curv), len(thick), len(elem), len(kap), len(ai), len(mat_names))
    return {
        "surfaces": [{"kind": kinds[i], "radius_mm": curv[i], "thickness_mm": thick[i],
                      "semi_diameter_mm": elem[i], "kappa": kap[i], "ai": ai[i]} for i in range(n)],
        "materials": [mats[x] for x in mat_names],
    }

def main():
    text = open(SRC).read()
    out = {}
    simple = {"tessar": "build_tessar_lens", "helios": "build_helios_lens", "jupiter": "build_jupiter_lens",
              "fisheye": "build_fisheye_lens", "gauss": "build_double_gauss_smith", "asph": "build_asph_lens"}
    for name, fn in simple.items():
        d = parse(body_of(text, fn))
        d.update(scale="1.0", stop_ratio=None, fine_focus=(name != "asph"))
        out[name] = d
    # exp1a/b/c: glass_B Cauchy pair passed at the call site (:883-888)
    for tag, (a, b) in {"exp1a": ("1.5689525390422485", "0.0263051608728981"),
                        "exp1b": ("1.507009332993809", "0.04781645616233479"),
                        "exp1c": ("1.5030211256389316", "0.04920145883751409")}.items():
        m = re.search(r'"' + tag + r'"\)\s*\{\s*build_exp1_doublet\(([^,]+),([^)]+)\)', text)
        assert num(m.group(1)) == a and num(m.group(2)) == b
        d = parse(body_of(text, "build_exp1_doublet"), subs={"mat_cauchy_A": a, "mat_cauchy_B": b})
        d.update(scale="1.0", stop_ratio=None, fine_focus=True)
        out[tag] = d
    # exp2a/b/c (:1452-1566): scale 35/1.57 (computed in float32 by the consumer), stop ratio 0.5
    body = body_of(text, "build_exp2_nikon")
    zero = ["0.0"] * 7
    m2 = re.search(r"mode == 2\).*?ai_list\.at\(16\)\s*=\s*\{([^}]*)\};\s*ai_list\.at\(17\)\s*=\s*\{([^}]*)\};", body, re.S)
    opt16 = [num(t) for t in m2.group(1).split(",")]; opt17 = [num(t) for t in m2.group(2).split(",")]
    for tag, ov in {"exp2a": None, "exp2b": {16: zero, 17: zero}, "exp2c": {16: opt16, 17: opt17}}.items():
        d = parse(body, ai_override=ov)
        d.update(scale="35.0/1.57", stop_ratio="0.5", fine_focus=True)
        out[tag] = d
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    for k, v in out.items():
        print(k, len(v["surfaces"]), "surfaces")

if __name__ == "__main__":
    main()
