"""The C ABI as Python sees it must be the C ABI as the header declares it: every `sig(...)` of
`kubeml_b200/_native/__init__.py` is compared with its prototype in `include/kubeml_b200.h` (argument count and the
class of every argument: pointer, integer width, float), and the ctypes structures with their C structs (field
count, order of names).  A drifted signature does not fail loudly on the GPU - it passes garbage."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _prototypes():
    text = open(os.path.join(ROOT, "include", "kubeml_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    protos = {}
    for m in re.finditer(r"\b([A-Za-z_][\w \*]*?)\b(kml_\w+)\s*\(([^;{}]*?)\)\s*;", text, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
        params = [] if args in ("", "void") else [a.strip() for a in args.split(",")]
        protos[name] = (ret, params)
    return protos


def _klass_of_c(param: str) -> str:
    p = param.replace("const ", "").strip()
    if "*" in p or "[" in p:
        return "ptr"
    t = " ".join(p.split()[:-1]) if len(p.split()) > 1 else p
    if t in ("float",):
        return "f32"
    if t in ("double",):
        return "f64"
    if t in ("size_t", "unsigned long long", "long long", "uint64_t", "int64_t"):
        return "i64"
    if t in ("int", "unsigned", "unsigned int", "uint32_t", "int32_t"):
        return "i32"
    raise AssertionError(f"unclassified C parameter: {param!r}")


def _klass_of_ctypes(t) -> str:
    if t in (ctypes.c_void_p, ctypes.c_char_p) or (isinstance(t, type) and issubclass(t, ctypes._Pointer)):
        return "ptr"
    if t is ctypes.c_float:
        return "f32"
    if t is ctypes.c_double:
        return "f64"
    if t in (ctypes.c_size_t, ctypes.c_ulonglong, ctypes.c_longlong, ctypes.c_uint64, ctypes.c_int64):
        return "i64"
    if t in (ctypes.c_int, ctypes.c_uint, ctypes.c_uint32, ctypes.c_int32):
        return "i32"
    raise AssertionError(f"unclassified ctypes type: {t!r}")


class _Recorder:
    """Stands in for the loaded library: `_declare` sets restype / argtypes on attributes of it."""

    def __init__(self):
        self.fns = {}

    def __getattr__(self, name):
        if name.startswith("kml_"):
            return self.fns.setdefault(name, type("F", (), {})())
        raise AttributeError(name)


def test_every_declared_signature_matches_the_header():
    from kubeml_b200 import _native
    rec = _Recorder()
    _native._declare(rec)
    protos = _prototypes()
    assert len(rec.fns) >= 60 and len(protos) >= len(rec.fns)
    for name, fn in rec.fns.items():
        assert name in protos, f"{name} is bound in Python but not declared in include/kubeml_b200.h"
        ret, params = protos[name]
        assert len(fn.argtypes) == len(params), f"{name}: {len(fn.argtypes)} ctypes arguments, {len(params)} in the header"
        for i, (ct, cp) in enumerate(zip(fn.argtypes, params)):
            assert _klass_of_ctypes(ct) == _klass_of_c(cp), f"{name}: argument {i} is {ct} in Python, `{cp}` in the header"
        if fn.restype is None:
            assert ret == "void", name
        elif fn.restype is ctypes.c_char_p:
            assert "char" in ret and "*" in ret, name
        else:
            assert _klass_of_ctypes(fn.restype) == _klass_of_c(ret + " x"), f"{name}: return type"


def _struct_fields(text: str, cname: str):
    """Field names of `typedef struct { ... } cname;` in declaration order (several declarators per statement)."""
    m = re.search(r"typedef\s+struct\s+\w*\s*\{([^{}]*)\}\s*" + cname + r"\s*;", text, flags=re.S)
    assert m, f"struct {cname} not found in the header"
    names = []
    for decl in m.group(1).split(";"):
        decl = " ".join(decl.split())
        if not decl:
            continue
        for i, part in enumerate(decl.split(",")):
            tok = part.split()[-1] if i == 0 else part.strip()
            names.append(re.sub(r"\[.*?\]", "", tok).lstrip("*").strip())
    return names


def test_ctypes_structures_mirror_the_c_structs():
    from kubeml_b200 import _native
    text = open(os.path.join(ROOT, "include", "kubeml_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    # the kernels read these by offset: same fields, same order
    assert _struct_fields(text, "kml_merge_params") == [f[0] for f in _native.MergeParams._fields_]
    assert _struct_fields(text, "kml_bn_fuse") == [f[0] for f in _native.BnFuse._fields_]
    assert _struct_fields(text, "kml_epilogue") == [f[0] for f in _native.Epilogue._fields_]
