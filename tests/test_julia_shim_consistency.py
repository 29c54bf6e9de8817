"""julia/DCEGMCuda.jl cannot be executed here (no julia).  What CAN be checked mechanically: its POD mirrors list the
fields of include/dcegm.h's structs in order and with the right types, every `ccall` names a function the header
declares, with the declared number of arguments, and the status bits it tests carry the header's values."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
JL = open(os.path.join(ROOT, "julia", "DCEGMCuda.jl")).read()
HDR = open(os.path.join(ROOT, "include", "dcegm.h")).read()
HDR_NC = re.sub(r"/\*.*?\*/", "", HDR, flags=re.S)

CTYPE = {"double": "Float64", "int32_t": "Int32", "int64_t": "Int64"}


def c_struct_fields(name):
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), HDR_NC, re.S).group(1)
    out = []
    for decl in body.split(";"):
        decl = " ".join(decl.split())
        if not decl:
            continue
        m = re.match(r"(const )?(\w+)\s*(\*?)\s*(.*)", decl)
        typ, ptr, names = m.group(2), m.group(3), m.group(4)
        for n in names.split(","):
            n = n.strip()
            p = ptr or ("*" if n.startswith("*") else "")
            out.append((n.lstrip("* "), ("Ptr{%s}" % CTYPE[typ]) if p else CTYPE[typ]))
    return out


def jl_struct_fields(name):
    body = re.search(r"struct %s\b.*?\n(.*?)\nend" % name, JL, re.S).group(1)
    out = []
    for f in re.findall(r"(\w+)::([\w{}]+)", body):
        out.append(f)
    return out


def test_param_mirror_matches_header():
    c = c_struct_fields("dcegm_param")
    j = jl_struct_fields("CParam")
    assert [n.replace("lambda", "lambda") for n, _ in c] == [n for n, _ in j]
    assert [t for _, t in c] == [t for _, t in j]
    assert len(c) == 23 + 8 + 1


def test_layout_and_shock_mirrors_match_header():
    assert c_struct_fields("dcegm_layout") == jl_struct_fields("CLayout")
    assert c_struct_fields("dcegm_sim_shocks") == jl_struct_fields("CSimShocks")


def header_prototypes():
    protos = {}
    for m in re.finditer(r"\b(dcegm_[a-z_0-9]+)\s*\(([^;{]*?)\)\s*;", HDR_NC, re.S):
        args = " ".join(m.group(2).split())
        protos[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    return protos


def test_every_ccall_names_a_declared_function_with_the_right_arity():
    protos = header_prototypes()
    calls = re.findall(r"ccall\(\(\s*:?(\w+)\s*,\s*LIB\)\s*,\s*\w+\s*,\s*\((.*?)\)\s*,", JL, re.S)
    assert len(calls) >= 15
    for sym, argt in calls:
        if sym in ("sym", "f"):  # symbol chosen at run time from names checked below
            continue
        assert sym in protos, f"{sym} is not declared in include/dcegm.h"
        depth, n, cur = 0, 0, ""
        for ch in argt:  # count top-level commas of the Julia argument-type tuple
            depth += ch in "{(" 
            depth -= ch in "})"
            if ch == "," and depth == 0:
                n += 1 if cur.strip() else 0
                cur = ""
            else:
                cur += ch
        n += 1 if cur.strip() else 0
        assert n == protos[sym], f"{sym}: {n} argument types in the shim, {protos[sym]} in the header"
    for sym in re.findall(r":(dcegm_\w+)", JL):
        assert sym in protos, sym


def test_status_bits_in_the_shim_carry_the_header_values():
    bits = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define DCEGM_ST_(\w+)\s+\(1u << (\d+)\)", HDR)}
    used = {0x0004: "NONMONOTONE", 0x4000: "DUP_KNOT", 0x0008: "NO_BRACKET", 0x0010: "MISALIGNED", 0x0020: "ALL_FALLING",
            0x0400: "NEG_CONS"}
    for val, name in used.items():
        assert 1 << bits[name] == val and ("0x%04x" % val) in JL
    assert (1 << bits["EXTRA_RIGHT"]) | (1 << bits["NONMONO_FIRST"]) == 0x1000 | 0x2000
    assert "const ST_INFO_MASK = 0x1000 | 0x2000" in JL
