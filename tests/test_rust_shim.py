"""The Rust side of the boundary (rust/ffi.rs, rust/build.rs) cannot be compiled on this image — no rustc, no cargo — so it is
held against include/frb.h and csrc/Makefile as TEXT: every `extern "C"` prototype (argument count, integer widths, pointer
constness), every #[repr(C)] struct (field order and types), the list of translation units and the nvcc flags.  A drifted
binding is a silent ABI break for the one consumer north_star names (the reference's get_image, src/lib.rs:253)."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

C_BASE = {"uint8_t": "u8", "uint32_t": "u32", "int32_t": "i32", "uint64_t": "u64", "int64_t": "i64", "size_t": "usize", "int": "c_int",
          "double": "f64", "float": "f32", "char": "c_char", "void": "c_void"}


def _read(*parts):
    return open(os.path.join(ROOT, *parts)).read()


def _header():
    src = re.sub(r"/\*.*?\*/", "", _read("include", "frb.h"), flags=re.S)
    return re.sub(r"//[^\n]*", "", src)


def c_type_to_rust(t: str) -> str:
    """`const uint8_t **` -> `*mut *const u8`: const binds to the innermost pointee, outer pointers are mutable."""
    t = t.strip()
    stars = t.count("*")
    toks = t.replace("*", " ").split()
    const = "const" in toks
    base = [k for k in toks if k not in ("const", "struct")]
    assert len(base) == 1, t
    r = C_BASE.get(base[0], base[0])
    for i in range(stars):
        r = ("*const " if (const and i == 0) else "*mut ") + r
    return r


def c_prototypes():
    out = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(frb_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", _header()):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        params = []
        if args and args != "void":
            for a in args.split(","):
                mm = re.match(r"(.*?)([A-Za-z_]\w*)$", a.strip())
                params.append(c_type_to_rust(mm.group(1)))
        out[name] = (params, None if ret == "void" else c_type_to_rust(ret))
    return out


def rust_prototypes():
    src = re.sub(r"//[^\n]*", "", _read("rust", "ffi.rs"))
    block = re.search(r'extern "C" \{(.*?)\n\}', src, flags=re.S).group(1)
    out = {}
    for m in re.finditer(r"pub fn (frb_[a-z0-9_]+)\s*\((.*?)\)\s*(?:->\s*([^;]+?))?\s*;", block, flags=re.S):
        params = [" ".join(a.split(":", 1)[1].split()) for a in m.group(2).split(",") if a.strip()]
        out[m.group(1)] = (params, " ".join(m.group(3).split()) if m.group(3) else None)
    return out


def c_structs():
    out = {}
    for m in re.finditer(r"typedef struct (\w+) \{(.*?)\} \1;", _header(), flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if decl:
                mm = re.match(r"(.*?)([A-Za-z_]\w*)$", decl)
                fields.append((mm.group(2), c_type_to_rust(mm.group(1))))
        out[m.group(1)] = fields
    return out


def rust_structs():
    src = re.sub(r"//[^\n]*", "", _read("rust", "ffi.rs"))
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\]\s*(?:#\[derive\([^)]*\)\]\s*)?pub struct (\w+) \{(.*?)\}", src, flags=re.S):
        out[m.group(1)] = [(f.split(":")[0].replace("pub", "").strip(), f.split(":")[1].strip()) for f in m.group(2).split(",") if f.strip()]
    return out


def test_every_header_function_is_bound_with_the_same_signature():
    c, r = c_prototypes(), rust_prototypes()
    assert len(c) >= 42
    assert sorted(c) == sorted(r), (sorted(set(c) - set(r)), sorted(set(r) - set(c)))
    for name in c:
        assert r[name] == c[name], f"{name}: header {c[name]} vs ffi.rs {r[name]}"


def test_repr_c_structs_mirror_the_header_field_for_field():
    c, r = c_structs(), rust_structs()
    assert {"FrbImaginary", "FrbRGB", "FrbConfig", "FrbShard", "FrbStats", "FrbServerStats", "FrbFp64Peak"} <= set(c)
    for name, fields in c.items():
        assert name in r, f"ffi.rs has no #[repr(C)] struct {name}"
        assert r[name] == fields, f"{name}: header {fields} vs ffi.rs {r[name]}"
    # the layout the header's static asserts pin, recomputed from the Rust field list (natural alignment, as #[repr(C)] lays out)
    size = {"u8": 1, "u32": 4, "f32": 4, "f64": 8, "u64": 8, "FrbImaginary": 16, "FrbRGB": 3}
    align = {"u8": 1, "u32": 4, "f32": 4, "f64": 8, "u64": 8, "FrbImaginary": 8, "FrbRGB": 1}
    off, offsets = 0, {}
    for fname, ftype in r["FrbConfig"]:
        off = (off + align[ftype] - 1) // align[ftype] * align[ftype]
        offsets[fname] = off
        off += size[ftype]
    assert (off + 7) // 8 * 8 == 104
    assert [offsets[k] for k in ("algo", "width", "height", "iterations", "limit", "stable_limit", "pos", "scale", "exposure", "inside",
                                 "smooth", "primary_color", "secondary_color", "color_weight", "julia_set")] == \
        [0, 4, 8, 12, 16, 24, 32, 48, 64, 72, 73, 74, 77, 80, 88]


def test_build_rs_compiles_what_the_makefile_compiles_with_the_same_flags():
    mk = _read("fractal_renderer_b200", "csrc", "Makefile")
    src = re.search(r"^SRC\s*:=\s*(.*)$", mk, flags=re.M).group(1).split()
    build = _read("rust", "build.rs")
    listed = re.findall(r'"(frb_\w+\.cu)"', build)
    assert sorted(listed) == sorted(src), f"build.rs {listed} vs Makefile {src}"
    for f in listed:
        assert os.path.exists(os.path.join(ROOT, "fractal_renderer_b200", "csrc", f)), f
    for flag in ("arch=compute_100a,code=sm_100a", "-fmad=false", "-lineinfo", "-std=c++17", "-shared", "-fPIC"):
        assert flag in build, flag
        assert flag in mk, flag


def test_the_drop_in_uses_the_bound_calls_the_reference_callers_need():
    """get_image (src/lib.rs:253) -> frb_render / frb_render_fern; the GUI thread (src/gui.rs:56-82) -> frb_server_*."""
    src = _read("rust", "ffi.rs")
    for call in ("frb_render(&cfg", "frb_render_fern(&cfg", "frb_server_create(", "frb_server_request(", "frb_server_acquire(", "frb_server_destroy(",
                 "frb_host_alloc(", "frb_host_free("):
        assert call in src, call
