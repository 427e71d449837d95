"""rust/src/ffi.rs is written by hand (no bindgen, no Rust toolchain in this image: the crate under rust/ is UNCOMPILED).  What
can be checked without rustc is that its `extern "C"` block has not drifted from include/linalg_b200.h: every function it
declares exists in the header with the same number of parameters, the same kind of each parameter (integer / size / pointer /
callback) and the same kind of result."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def split_params(text):
    out, depth, cur = [], 0, ""
    for ch in text:
        if ch in "(<":
            depth += 1
        elif ch in ")>":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip()); cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def c_kind(param):
    p = param.strip()
    if p == "void":
        return None
    if "*" in p:
        return "ptr"
    if p.startswith("la_exchange_fn"):
        return "fn"
    if re.match(r"(const\s+)?size_t\b", p):
        return "size"
    if re.match(r"(const\s+)?uint64_t\b", p):
        return "u64"
    if re.match(r"(const\s+)?int\b", p):
        return "int"
    raise AssertionError("unclassified C parameter: %r" % param)


def rust_kind(param):
    ty = param.split(":", 1)[1].strip()
    if ty.startswith("*") or ty == "Stream":
        return "ptr"
    if ty.startswith("Option<"):
        return "fn"
    return {"usize": "size", "c_int": "int", "u64": "u64"}[ty]


def header_functions():
    src = open(os.path.join(ROOT, "include", "linalg_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    fns = {}
    for m in re.finditer(r"^\s*([A-Za-z_][\w\s\*]*?)\b(la_[a-z0-9_]+)\s*\(([^;{]*)\)\s*;", src, flags=re.M):
        ret, name, params = m.group(1).strip(), m.group(2), m.group(3)
        if ret.startswith("typedef"):
            continue
        kinds = [k for k in (c_kind(p) for p in split_params(params)) if k]
        fns[name] = ("void" if ret == "void" else "ptr" if "*" in ret else "size" if ret == "size_t" else "int", kinds)
    return fns


def rust_functions():
    src = open(os.path.join(ROOT, "rust", "src", "ffi.rs")).read()
    block = src[src.index('extern "C" {'):]
    block = block[:block.index("\n}\n")]
    fns = {}
    for m in re.finditer(r"pub fn (la_[a-z0-9_]+)\((.*?)\)\s*(?:->\s*([^;]+))?;", block, flags=re.S):
        name, params, ret = m.group(1), m.group(2), (m.group(3) or "").strip()
        kinds = [rust_kind(p) for p in split_params(params)]
        fns[name] = ("void" if not ret else "ptr" if ret.startswith("*") else "size" if ret == "usize" else "int", kinds)
    return fns


def test_rust_ffi_matches_the_header():
    c, r = header_functions(), rust_functions()
    assert len(r) >= 35 and len(c) >= 60
    for name, (ret, kinds) in r.items():
        assert name in c, "rust/src/ffi.rs declares %s, which include/linalg_b200.h does not" % name
        assert c[name][1] == kinds, "%s: header takes %s, ffi.rs declares %s" % (name, c[name][1], kinds)
        assert c[name][0] == ret, "%s: header returns %s, ffi.rs declares %s" % (name, c[name][0], ret)


def test_the_facade_can_reach_the_whole_hot_path():
    """Every kernel family of src/kernels.cl plus transpose, the prepared-B product and the multi-GPU entry points."""
    r = rust_functions()
    for name in ("la_ew_vec_vec", "la_ew_assign_vec_vec", "la_ew_vec_scl", "la_ew_assign_vec_scl", "la_eq_vec", "la_sum_vec", "la_dot_vec_vec",
                 "la_mul_vec_mat", "la_mul_vec_transpose_mat", "la_mul_mat_mat", "la_transpose", "la_gemm_prepare_b", "la_mul_mat_mat_prepared",
                 "la_comm_create", "la_sum_vec_sharded", "la_dot_vec_vec_sharded", "la_mul_vec_mat_sharded", "la_transpose_sharded", "la_mul_mat_mat_host_sharded"):
        assert name in r, name
