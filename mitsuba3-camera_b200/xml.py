"""Scene-description XML -> the nested dictionaries `load_dict` takes: the subset of src/core/xml.cpp the fork's own scenes
(notebooks/scenes/**/*.xml) use — `<default>` + `$name` substitution (:402-420, longest name first), typed properties
(`float`, `integer`, `boolean`, `string`, `point`, `vector`, `rgb`, constant `spectrum`), `<transform>` built from `translate`,
`rotate`, `scale`, `lookat`, `matrix` (each new operation multiplies from the LEFT, :861-931), `<ref>`, `<include>`, nested plugin
objects. `mi.load_file(path, **defaults)` / `mi.load_string(text, **defaults)` as in src/core/python/xml_v.cpp.

Naming follows xml.cpp:486-504: an object without a `name` attribute becomes `_arg_<i>` (i counts the unnamed objects and
references of its parent), its `id` is kept under 'id'. The reference's property map orders such names numerically
(scene.natural_key), so a scene's shapes and emitters keep their document order. Unnamed `film`, `sampler`, `rfilter` children
take their tag as key — the names the sensor / film dictionaries of `load_dict` use."""
from __future__ import annotations

import os
import xml.etree.ElementTree as ET

import numpy as np

from .transform import Transform4f

f32 = np.float32
_OBJECT_TAGS = ("scene", "shape", "bsdf", "emitter", "sensor", "film", "sampler", "rfilter", "integrator", "texture", "medium",
                "phase", "volume", "spectrum_object")
_KEY_BY_TAG = ("film", "sampler", "rfilter")


class XMLError(ValueError):
    pass


def _floats(text, what):
    try:
        return [float(t) for t in text.replace(",", " ").split()]
    except ValueError as e:
        raise XMLError(f"could not parse floating point values in {what} \"{text}\"") from e


def _vec3(el, default=0.0, what="vector"):
    """detail::expand_value_to_xyz + parse_vector (xml.cpp): value=\"a\" | \"a, b, c\" or x= y= z= (missing -> default)."""
    if "value" in el.attrib:
        v = _floats(el.attrib["value"], what)
        if len(v) == 1:
            return [v[0]] * 3
        if len(v) != 3:
            raise XMLError(f"'value' attribute of <{el.tag}> must have exactly 1 or 3 elements")
        return v
    return [float(el.attrib.get(a, default)) for a in "xyz"]


def _named_vec(el, name):
    v = _floats(el.attrib[name], name)
    if len(v) != 3:
        raise XMLError(f"<{el.tag}>: '{name}' must have three components")
    return v


def _transform(el):
    T = Transform4f()
    for op in el:
        if op.tag == "translate":
            M = Transform4f.translate(_vec3(op))
        elif op.tag == "scale":
            M = Transform4f.scale(_vec3(op, 1.0))
        elif op.tag == "rotate":
            M = Transform4f.rotate([float(op.attrib.get(a, 0.0)) for a in "xyz"], float(op.attrib["angle"]))
        elif op.tag == "lookat":
            origin, target = _named_vec(op, "origin"), _named_vec(op, "target")
            up = _named_vec(op, "up") if "up" in op.attrib else [0.0, 0.0, 0.0]
            if not any(up):     # xml.cpp:887-889: any vector perpendicular to the viewing direction
                d = np.asarray(target, np.float64) - np.asarray(origin, np.float64); d /= np.linalg.norm(d)
                s = np.copysign(1.0, d[2]); a = -1.0 / (s + d[2]); b = d[0] * d[1] * a
                up = [s * d[0] * d[0] * a + 1.0, s * b, -s * d[0]]
            M = Transform4f.look_at(origin=origin, target=target, up=up)
        elif op.tag == "matrix":
            v = _floats(op.attrib["value"], "matrix")
            if len(v) == 16:
                M = Transform4f(np.array(v, f32).reshape(4, 4))
            elif len(v) == 9:
                m = np.eye(4, dtype=f32); m[:3, :3] = np.array(v, f32).reshape(3, 3); M = Transform4f(m)
            else:
                raise XMLError("matrix: expected 16 or 9 values")
        else:
            raise XMLError(f"transform nodes can only contain transform operations (found <{op.tag}>)")
        T = M @ T
    return T


class _Loader:
    def __init__(self, params):
        self.params = {str(k): str(v) for k, v in params.items()}
        self.used = set()

    def subst(self, el):
        """`$name` in every attribute (longest names first); an undefined parameter is an error (xml.cpp:402-420)."""
        for k, v in list(el.attrib.items()):
            if "$" not in v:
                continue
            for name in sorted(self.params, key=len, reverse=True):
                if "$" + name in v:
                    v = v.replace("$" + name, self.params[name]); self.used.add(name)
            if "$" in v:
                raise XMLError(f"undefined parameter(s) in string: \"{v}\"!")
            el.attrib[k] = v

    def parse_file(self, path):
        try:
            root = ET.parse(path).getroot()
        except ET.ParseError as e:
            raise XMLError(f"Error while loading \"{path}\": {e}") from e
        return root

    def object(self, el, base_dir, is_root=False):
        self.subst(el)
        if is_root and "version" not in el.attrib:
            raise XMLError(f"missing version attribute in root element \"{el.tag}\"")
        d = {"type": "scene" if el.tag == "scene" else el.attrib.get("type")}
        if d["type"] is None:
            raise XMLError(f"missing attribute \"type\" in element \"{el.tag}\"")
        if "id" in el.attrib:
            d["id"] = el.attrib["id"]
        counter = [0]
        self.children(el, d, base_dir, counter, scene=el.tag == "scene")
        return d

    def children(self, el, d, base_dir, counter, scene):
        def put(key, value, node):
            if key in d:
                raise XMLError(f"<{node.tag}>: \"{key}\" was already specified")
            d[key] = value

        def auto_name(node):
            i = counter[0]; counter[0] += 1
            if not scene and node.tag in _KEY_BY_TAG and node.tag not in d:
                return node.tag
            return f"_arg_{i}"
        for ch in el:
            if not isinstance(ch.tag, str):      # comments / processing instructions
                continue
            self.subst(ch)
            tag, name = ch.tag, ch.attrib.get("name")
            if name is not None and name.startswith("_"):
                raise XMLError(f"invalid parameter name \"{name}\" in element \"{tag}\": leading underscores are reserved for internal identifiers.")
            if tag == "default":
                n = ch.attrib["name"]
                if n not in self.params:
                    self.params[n] = ch.attrib["value"]
            elif tag in ("alias", "path"):
                continue
            elif tag == "include":
                fn = ch.attrib["filename"]
                fn = fn if os.path.isabs(fn) else os.path.join(base_dir, fn)
                root = self.parse_file(fn)
                self.subst(root)
                if root.tag == "scene":
                    self.children(root, d, os.path.dirname(os.path.abspath(fn)), counter, scene)
                else:
                    put(root.attrib.get("name") or auto_name(root), self.object(root, os.path.dirname(os.path.abspath(fn))), root)
            elif tag in _OBJECT_TAGS or (tag == "spectrum" and "type" in ch.attrib):
                put(name or auto_name(ch), self.object(ch, base_dir), ch)
            elif tag == "ref":
                put(name or auto_name(ch), {"type": "ref", "id": ch.attrib["id"]}, ch)
            elif name is None:
                raise XMLError(f"<{tag}>: missing attribute \"name\"")
            elif tag == "float":
                put(name, float(ch.attrib["value"]), ch)
            elif tag == "integer":
                try:
                    put(name, int(ch.attrib["value"]), ch)
                except ValueError as e:
                    raise XMLError(f"could not parse integer value \"{ch.attrib['value']}\"") from e
            elif tag == "boolean":
                v = ch.attrib["value"].lower()
                if v not in ("true", "false"):
                    raise XMLError(f"could not parse boolean value \"{v}\" -- must be \"true\" or \"false\"")
                put(name, v == "true", ch)
            elif tag == "string":
                v = ch.attrib["value"]
                if name == "filename" and not os.path.isabs(v):
                    v = os.path.join(base_dir, v)
                put(name, v, ch)
            elif tag in ("point", "vector"):
                put(name, _vec3(ch, 0.0, tag), ch)
            elif tag == "rgb":
                v = _floats(ch.attrib["value"], "rgb")
                if len(v) == 1:
                    v = v * 3
                if len(v) != 3:
                    raise XMLError("'rgb' tag requires one or three values (got \"%s\")" % ch.attrib["value"])
                put(name, {"type": "rgb", "value": v}, ch)
            elif tag == "spectrum":
                if "filename" in ch.attrib or ":" in ch.attrib.get("value", ""):
                    raise XMLError("<spectrum>: only a constant value is built (tabulated spectra belong to the spectral variants)")
                v = _floats(ch.attrib["value"], "spectrum")
                if len(v) != 1:
                    raise XMLError("<spectrum>: expected a single value")
                put(name, {"type": "spectrum", "value": v[0]}, ch)
            elif tag == "transform":
                put(name, _transform(ch), ch)
            else:
                raise XMLError(f"unexpected tag \"{tag}\"")


def xml_to_dict(path=None, text=None, **params):
    """The dictionary `load_dict` would be given for this scene description (file path or XML text)."""
    ld = _Loader(params)
    if path is not None:
        root, base = ld.parse_file(path), os.path.dirname(os.path.abspath(path))
    else:
        try:
            root = ET.fromstring(text)
        except ET.ParseError as e:
            raise XMLError(f"Error while loading XML string: {e}") from e
        base = os.getcwd()
    if root.tag not in _OBJECT_TAGS:
        raise XMLError(f"root element \"{root.tag}\" must be an object")
    d = ld.object(root, base, is_root=True)
    unused = [k for k in params if str(k) not in ld.used]
    if unused:       # xml.cpp: "Unused parameter(s)" is an error there too
        raise XMLError("Found unused parameters: " + ", ".join(map(str, unused)))
    return d


def load_file(path, **params):
    from .scene import load_dict
    if not os.path.isfile(path):
        raise XMLError(f"\"{path}\": file does not exist!")
    return load_dict(xml_to_dict(path=path, **params))


def load_string(text, **params):
    from .scene import load_dict
    return load_dict(xml_to_dict(text=text, **params))
