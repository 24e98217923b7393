#!/usr/bin/env python3
"""Extract the D4 parameter DATA (not code) from the reference tree into one C include file.

Reads (read-only):
  src/dftd4/reference.inc            reference systems: refn, refsys, clsq/clsh, hcount, ascale,
                                     refcovcn, refcn, alphaiw(23,7,118), sscale, secaiw(23,17)
  src/dftd4/data/{covrad,en,hardness,r4r2,zeff}.f90   per-element scalars
  src/dftd4/param.f90                functional name -> id, BJ and BJ-ATM damping parameter tables

Writes dftd4_b200/csrc/d4_tables_data.inc: plain C arrays, indexed by atomic number Z (slot 0 unused).
Numbers are carried as the shortest decimal string that round-trips the double the Fortran literal denotes,
so the generated tables are bit-identical to what a Fortran compiler stores.

The generated file is committed (the GPU box has no /root/reference); re-run only when the reference changes.
"""
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = sys.argv[2] if len(sys.argv) > 2 else os.path.join(
    os.path.dirname(os.path.abspath(__file__)), "..", "dftd4_b200", "csrc", "d4_tables_data.inc")
SRC = os.path.join(REF, "src", "dftd4")
MAXZ = 118
MREF = 7
NFREQ = 23
NSEC = 17


def fnum(tok):
    tok = tok.strip().replace("_wp", "").replace("d", "e").replace("D", "e")
    return float(tok)


def statements(path):
    """Fortran fixed/free mixed include file -> list of logical statements (continuations joined)."""
    out, cur = [], ""
    for raw in open(path, encoding="utf-8"):
        line = raw.rstrip("\n")
        if line.lstrip().startswith("!"):
            continue
        if "!" in line:
            line = line[:line.index("!")]
        s = line.strip()
        if not s:
            continue
        if s.startswith("&"):
            cur += " " + s[1:].strip()
        else:
            if cur:
                out.append(cur)
            cur = s
        if cur.endswith("&"):
            cur = cur[:-1].strip()
    if cur:
        out.append(cur)
    res = []
    for st in out:
        res.extend(x.strip() for x in st.split(";") if x.strip())
    return res


def parse_reference_inc():
    d = {
        "refn": [0] * (MAXZ + 1),
        "refsys": [[0] * MREF for _ in range(MAXZ + 1)],
        "alphaiw": [[[0.0] * NFREQ for _ in range(MREF)] for _ in range(MAXZ + 1)],
        "sscale": [0.0] * (NSEC + 1),
        "secaiw": [[0.0] * NFREQ for _ in range(NSEC + 1)],
    }
    for k in ("clsq", "clsh", "hcount", "ascale", "refcovcn", "refcn"):
        d[k] = [[0.0] * MREF for _ in range(MAXZ + 1)]
    pat = re.compile(r"^data\s+(\w+)\s*\(([^)]*)\)\s*/(.*)/\s*$", re.I)
    for st in statements(os.path.join(SRC, "reference.inc")):
        m = pat.match(st)
        if not m:
            raise SystemExit("unparsed statement: " + st[:80])
        name, idx, vals = m.group(1).lower(), [x.strip() for x in m.group(2).split(",")], m.group(3)
        toks = [t for t in vals.split(",") if t.strip()]
        if name == "refn":
            d["refn"][int(idx[0])] = int(toks[0])
        elif name == "refsys":
            d["refsys"][int(idx[1])][int(idx[0]) - 1] = int(toks[0])
        elif name in ("clsq", "clsh", "hcount", "ascale", "refcovcn", "refcn"):
            d[name][int(idx[1])][int(idx[0]) - 1] = fnum(toks[0])
        elif name == "alphaiw":
            assert idx[0] == ":" and len(toks) == NFREQ
            d["alphaiw"][int(idx[2])][int(idx[1]) - 1] = [fnum(t) for t in toks]
        elif name == "sscale":
            d["sscale"][int(idx[0])] = fnum(toks[0])
        elif name == "secaiw":
            assert idx[0] == ":" and len(toks) == NFREQ
            d["secaiw"][int(idx[1])] = [fnum(t) for t in toks]
        # every other array (refq, gffq, dftq, pbcq, eeqbcq, secq, seccn, ...) belongs to charge models
        # outside the hot path (GFN2 / EEQ-BC) and is not extracted.
    return d


def parse_array(path, name):
    """Pull `real(wp), parameter :: name(max_elem) = [factor *] [ ... ]` out of a data module."""
    txt = open(path, encoding="utf-8").read()
    m = re.search(r"::\s*" + name + r"\s*\(max_elem\)\s*=\s*([^\[]*)\[(.*?)\]", txt, re.S)
    body = re.sub(r"!.*", "", m.group(2))
    body = body.replace("&", " ")
    toks = [t for t in re.split(r"[,\s]+", body) if t]
    vals = [fnum(t) for t in toks]
    assert len(vals) == MAXZ, (name, len(vals))
    return [0.0] + vals


def parse_param():
    txt = open(os.path.join(SRC, "param.f90"), encoding="utf-8").read()
    lines = txt.split("\n")
    # enumerators
    m = re.search(r"enumerator\s*::\s*p_invalid\s*,(.*?)end enum", txt, re.S)
    names = [t.strip() for t in re.sub(r"[&\n]", " ", m.group(1)).split(",") if t.strip()]
    ids = {"p_invalid": 0}
    for i, n in enumerate(names):
        ids[n] = i + 1

    def table(start_pat, end_pat):
        s = next(i for i, l in enumerate(lines) if re.match(start_pat, l))
        e = next(i for i, l in enumerate(lines) if re.match(end_pat, l))
        body = "\n".join(re.sub(r"!.*", "", l) for l in lines[s:e])
        body = body.split("contains")[0]
        out = {}
        for cm in re.finditer(r"case\((p_\w+)\)\s*param\s*=\s*dftd_param\s*\((.*?)\)\s*(?=case|end select)", body, re.S):
            args = re.sub(r"[&\n]", " ", cm.group(2))
            kv = dict((k.strip(), fnum(v)) for k, v in (a.split("=") for a in args.split(",") if "=" in a))
            out[cm.group(1)] = (kv.get("s6", 1.0), kv["s8"], kv["a1"], kv["a2"], kv.get("alp", 16.0))
        return out

    bj = table(r"^subroutine get_d4eeq_bj_parameter", r"^end subroutine get_d4eeq_bj_parameter")
    bjatm = table(r"^subroutine get_d4eeq_bjatm_parameter", r"^end subroutine get_d4eeq_bjatm_parameter")
    s = next(i for i, l in enumerate(lines) if re.match(r"^pure function get_functional_id", l))
    e = next(i for i, l in enumerate(lines) if re.match(r"^end function get_functional_id", l))
    body = "\n".join(re.sub(r'!(?=[^"]*$).*', "", l) for l in lines[s:e])
    alias = []
    for cm in re.finditer(r"case\(((?:\s*\"[^\"]*\"\s*,?\s*&?\s*)+)\)\s*num\s*=\s*(p_\w+)", body, re.S):
        for nm in re.findall(r"\"([^\"]*)\"", cm.group(1)):
            alias.append((nm, cm.group(2)))
    return ids, bj, bjatm, alias


def parse_wfpair():
    """data/wfpair.f90: `data weight_factors(:, j) / ... /` columns of the D4S pair weighting table (118 x 118);
    get_wfpair_val(zi, zj) = weight_factors(zi, zj) (wfpair.f90:3836-3849).  Returned as tab[zi][zj], 1-based."""
    tab = [[0.0] * (MAXZ + 1) for _ in range(MAXZ + 1)]
    pat = re.compile(r"^data\s+weight_factors\s*\(\s*:\s*,\s*(\d+)\s*\)\s*/(.*)/\s*$", re.I)
    seen = 0
    for st in statements(os.path.join(SRC, "data", "wfpair.f90")):
        m = pat.match(st)
        if not m:
            continue
        j = int(m.group(1))
        vals = [fnum(t) for t in m.group(2).split(",") if t.strip()]
        assert len(vals) == MAXZ, (j, len(vals))
        for i, v in enumerate(vals):
            tab[i + 1][j] = v
        seen += 1
    assert seen == MAXZ, seen
    return tab


def cfloat(x):
    r = repr(float(x))
    if "e" not in r and "." not in r and "inf" not in r and "nan" not in r:
        r += ".0"
    return r


def emit_1d(f, ctype, name, vals, per=6):
    f.write("static const %s %s[%d] = {\n" % (ctype, name, len(vals)))
    fmt = cfloat if ctype == "double" else str
    for i in range(0, len(vals), per):
        f.write("  " + ", ".join(fmt(v) for v in vals[i:i + per]) + ",\n")
    f.write("};\n\n")


def main():
    ref = parse_reference_inc()
    cov = parse_array(os.path.join(SRC, "data", "covrad.f90"), "covalent_rad_2009")
    en = parse_array(os.path.join(SRC, "data", "en.f90"), "pauling_en")
    hard = parse_array(os.path.join(SRC, "data", "hardness.f90"), "chemical_hardness")
    r4r2 = parse_array(os.path.join(SRC, "data", "r4r2.f90"), "r4_over_r2")
    zeff = parse_array(os.path.join(SRC, "data", "zeff.f90"), "effective_nuclear_charge")
    ids, bj, bjatm, alias = parse_param()

    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as f:
        f.write("// GENERATED by tools/gen_tables.py from the dftd4 v4.2.0 parameter data -- do not edit.\n")
        f.write("// Parameter DATA only (element constants, reference polarisabilities, damping parameters);\n")
        f.write("// index [Z] with Z = atomic number, slot 0 unused; reference index 0-based.\n")
        f.write("#define D4T_MAXZ %d\n#define D4T_MREF %d\n#define D4T_NFREQ %d\n#define D4T_NSEC %d\n\n" % (MAXZ, MREF, NFREQ, NSEC))
        emit_1d(f, "int", "d4t_refn", ref["refn"], 20)
        emit_1d(f, "int", "d4t_refsys", [v for row in ref["refsys"] for v in row], 21)
        for k in ("clsq", "clsh", "hcount", "ascale", "refcovcn", "refcn"):
            emit_1d(f, "double", "d4t_" + k, [v for row in ref[k] for v in row], 7)
        emit_1d(f, "double", "d4t_alphaiw", [v for z in ref["alphaiw"] for r in z for v in r], 6)
        emit_1d(f, "double", "d4t_sscale", ref["sscale"], 6)
        emit_1d(f, "double", "d4t_secaiw", [v for r in ref["secaiw"] for v in r], 6)
        emit_1d(f, "double", "d4t_covrad2009_angstrom", cov, 8)
        emit_1d(f, "double", "d4t_pauling_en", en, 8)
        emit_1d(f, "double", "d4t_hardness", hard, 6)
        emit_1d(f, "double", "d4t_r4_over_r2", r4r2, 8)
        emit_1d(f, "double", "d4t_zeff", zeff, 12)
        # damping parameters
        nid = max(ids.values()) + 1
        f.write("#define D4T_NFUNC %d\n" % nid)
        for nm, tab in (("bj", bj), ("bjatm", bjatm)):
            f.write("// {present, s6, s8, a1, a2, alp} by functional id\n")
            f.write("static const double d4t_param_%s[%d][6] = {\n" % (nm, nid))
            inv = {v: k for k, v in ids.items()}
            for i in range(nid):
                p = tab.get(inv[i])
                if p is None:
                    f.write("  {0,0,0,0,0,0},\n")
                else:
                    f.write("  {1, %s},  // %s\n" % (", ".join(cfloat(x) for x in p), inv[i][2:]))
            f.write("};\n\n")
        f.write("typedef struct { const char* name; int id; } d4t_alias_t;\n")
        f.write("static const d4t_alias_t d4t_alias[%d] = {\n" % len(alias))
        for nm, pid in alias:
            cname = "".join(c if 32 <= ord(c) < 127 and c not in '"\\' else "".join("\\%03o" % b for b in c.encode("utf-8")) for c in nm)
            f.write('  {"%s", %d},\n' % (cname, ids[pid]))
        f.write("};\n#define D4T_NALIAS %d\n" % len(alias))
    wf = parse_wfpair()
    out2 = os.path.join(os.path.dirname(OUT), "d4_wfpair_data.inc")
    with open(out2, "w") as f:
        f.write("// GENERATED by tools/gen_tables.py from src/dftd4/data/wfpair.f90 (dftd4 v4.2.0) -- do not edit.\n")
        f.write("// D4S pair-specific Gaussian weighting factors: d4t_wfpair[zi * 119 + zj] = weight_factors(zi, zj).\n")
        emit_1d(f, "double", "d4t_wfpair", [v for row in wf for v in row], 7)
    print("wrote", out2)
    print("wrote", OUT, "nfunc", nid, "bj", len(bj), "bjatm", len(bjatm), "alias", len(alias))


if __name__ == "__main__":
    main()
