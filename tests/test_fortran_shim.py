"""fortran/cfd_b200_shim.f90 against include/cfd_b200.h.  The build image has no Fortran compiler, so the shim cannot
be compiled here; this test parses it instead and checks everything a compiler + linker would trip over when the C
struct or an entry point changes: bind(c) types (component names, order, types, array extents), enum constants,
structure constructors (keyword form only, every keyword a component, every component default-initialised) and the
bind(c) interfaces (symbol exists in the header, same number of arguments, by-value where the C side is by value).
It also checks that the reference's dummy-argument lists are kept (lisolv / tdma) and that INTEGRATION.md shows the
same type definitions."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "fortran", "cfd_b200_shim.f90")
HEADER = os.path.join(ROOT, "include", "cfd_b200.h")


def _strip_fortran(src):
    out = []
    for ln in src.splitlines():
        if ln.lstrip().startswith("#"):
            continue
        q = ln.find("!")
        out.append(ln if q < 0 else ln[:q])
    text = "\n".join(out)
    return re.sub(r"&\s*\n\s*", " ", text)           # join continuation lines


def _strip_c(src):
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    return re.sub(r"//[^\n]*", " ", src)


def c_structs(hdr):
    """{name: [(component, ctype, extent or None), ...]} for every `typedef struct name { ... } name;`"""
    res = {}
    for m in re.finditer(r"typedef\s+struct\s+(\w+)\s*\{(.*?)\}\s*\1\s*;", hdr, flags=re.S):
        comps = []
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if not decl:
                continue
            ctype, rest = decl.split(None, 1)
            for item in rest.split(","):
                item = item.strip()
                am = re.match(r"(\w+)\s*\[\s*(\w+)\s*\]", item)
                if am:
                    comps.append((am.group(1), ctype, am.group(2)))
                else:
                    comps.append((item, ctype, None))
        res[m.group(1)] = comps
    return res


def c_enums(hdr):
    """All enumerators of the header -> value (C rules: previous + 1 unless given)."""
    vals = {}
    for m in re.finditer(r"enum\s*\w*\s*\{(.*?)\}", hdr, flags=re.S):
        cur = -1
        for item in m.group(1).split(","):
            item = item.strip()
            if not item:
                continue
            if "=" in item:
                name, v = [x.strip() for x in item.split("=")]
                cur = int(v, 0) if re.fullmatch(r"-?\w+", v) and not v[0].isalpha() else vals[v]
            else:
                name, cur = item, cur + 1
            vals[name] = cur
    return vals


def c_functions(hdr):
    """{name: [param strings]} for every prototype `int|const char*|long long name(...);`"""
    res = {}
    for m in re.finditer(r"(?:^|\n)\s*(?:int|long long|const char\s*\*)\s*(\w+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        params = [p.strip() for p in m.group(2).split(",")] if m.group(2).strip() not in ("", "void") else []
        res[m.group(1)] = params
    return res


def f_types(src):
    res = {}
    for m in re.finditer(r"type\s*,\s*bind\(c\)\s*::\s*(\w+)(.*?)end type", src, flags=re.S | re.I):
        comps = []
        for ln in m.group(2).splitlines():
            ln = ln.strip()
            if not ln:
                continue
            dm = re.match(r"(integer|real)\((c_int|c_double)\)\s*::\s*(.*)", ln, flags=re.I)
            assert dm, f"unparsed component line in type {m.group(1)}: {ln!r}"
            # split the entity list on commas outside parentheses
            ents, depth, cur = [], 0, ""
            for ch in dm.group(3):
                if ch == "(":
                    depth += 1
                if ch == ")":
                    depth -= 1
                if ch == "," and depth == 0:
                    ents.append(cur)
                    cur = ""
                else:
                    cur += ch
            ents.append(cur)
            for e in ents:
                em = re.match(r"\s*(\w+)\s*(?:\(\s*(\d+)\s*\))?\s*(=.*)?$", e)
                assert em, f"unparsed entity {e!r}"
                comps.append((em.group(1), dm.group(2), em.group(2), em.group(3) is not None))
        res[m.group(1)] = comps
    return res


def _load():
    shim = _strip_fortran(open(SHIM).read())
    hdr = _strip_c(open(HEADER).read())
    return shim, hdr


def test_bind_c_types_match_the_c_structs():
    shim, hdr = _load()
    cs, ft = c_structs(hdr), f_types(shim)
    enums = c_enums(hdr)
    assert {"cfd_config", "cfd_turbulence", "cfd_hook_desc", "cfd_heat_config"} <= set(ft)
    for name, fcomps in ft.items():
        ccomps = cs[name]
        assert [c[0] for c in fcomps] == [c[0] for c in ccomps], f"{name}: component names / order differ"
        for (fn, fk, fext, fdef), (cn, ctype, cext) in zip(fcomps, ccomps):
            assert {"int": "c_int", "double": "c_double"}[ctype] == fk, (name, fn)
            want = None if cext is None else str(enums.get(cext, cext))
            assert fext == want, (name, fn, fext, want)
            assert fdef, f"{name}%{fn} has no default initialisation (constructors rely on it)"


def test_structure_constructors_use_keywords_of_existing_components():
    shim, _ = _load()
    ft = f_types(shim)
    found = 0
    for tname, comps in ft.items():
        names = {c[0] for c in comps}
        for m in re.finditer(r"=\s*" + tname + r"\s*\(", shim):
            # balanced argument list
            i, depth = m.end(), 1
            while depth:
                depth += {"(": 1, ")": -1}.get(shim[i], 0)
                i += 1
            args, depth, cur = [], 0, ""
            for ch in shim[m.end():i - 1]:
                if ch in "([":
                    depth += 1
                if ch in ")]":
                    depth -= 1
                if ch == "," and depth == 0:
                    args.append(cur)
                    cur = ""
                else:
                    cur += ch
            if cur.strip():
                args.append(cur)
            found += 1
            for a in args:
                km = re.match(r"\s*(\w+)\s*=", a)
                assert km, f"positional argument {a.strip()!r} in a {tname} constructor: use keywords"
                assert km.group(1) in names, f"{tname} has no component {km.group(1)}"
    assert found >= 6         # cavity, its lisolv, general, tdma, TEACH, TEACH lisolv, hooks, turbulence


def test_constants_match_the_header_enums():
    shim, hdr = _load()
    enums = c_enums(hdr)
    n = 0
    for m in re.finditer(r"integer\(c_int\),\s*parameter\s*::(.*)", shim):
        for name, val in re.findall(r"(\w+)\s*=\s*(-?\d+)", m.group(1)):
            assert name in enums, name
            assert enums[name] == int(val), (name, enums[name], int(val))
            n += 1
    assert n >= 70
    for k in ("CFD_GAMH", "CFD_G_RCV", "CFD_WALL_XPLUSW", "CFD_HOOK_FIELD_SOURCES", "CFD_FAMILY_TEACH"):
        assert re.search(r"\b%s\s*=" % k, shim), k


def test_interfaces_exist_in_the_header_with_the_same_arity_and_value_passing():
    shim, hdr = _load()
    cf = c_functions(hdr)
    seen = 0
    for m in re.finditer(r"function\s+(\w+)\s*\(([^)]*)\)\s*bind\(c,\s*name='(\w+)'\)(.*?)end function", shim, flags=re.S):
        fname, fargs, cname, body = m.group(1), [a.strip() for a in m.group(2).split(",")], m.group(3), m.group(4)
        assert fname == cname and cname in cf, cname
        cparams = cf[cname]
        assert len(fargs) == len(cparams), (cname, fargs, cparams)
        by_value, handles = set(), set()
        for decl in re.split(r"[;\n]", body):
            if "::" not in decl:
                continue
            names = {re.sub(r"\(.*", "", x).strip() for x in re.split(r",(?![^()]*\))", decl.split("::")[1])}
            if re.search(r",\s*value\s*::", decl):
                by_value |= names
            if re.match(r"\s*type\(c_ptr\)", decl):
                handles |= names
        for fa, cp in zip(fargs, cparams):
            c_is_pointer = "*" in cp or "[" in cp
            if fa in handles:
                # an opaque handle: `cfd_ctx* ctx` <-> type(c_ptr), value; `cfd_ctx** out` <-> type(c_ptr) by reference
                assert (fa in by_value) == ("**" not in cp), (cname, fa, cp)
                assert "*" in cp, (cname, fa, cp)
                continue
            assert (fa in by_value) == (not c_is_pointer), (cname, fa, cp)
        seen += 1
    assert seen >= 35
    for need in ("cfd_lisolv", "cfd_calct", "cfd_upload_noalloc", "cfd_set_hook", "cfd_calcte", "cfd_calced", "cfd_props",
                 "cfd_heat_create", "cfd_heat_set_operator", "cfd_heat2d_steady_gs", "cfd_heat2d_implicit_step", "cfd_sipsol", "cfd_lin_solve"):
        assert re.search(r"name='%s'" % need, shim), need


def test_reference_subroutine_names_and_dummy_lists_are_kept():
    raw = open(SHIM).read().lower()
    # cavity: calc_mod + linsolv_mod (cavity_fvm.f90:43-82, 84-357)
    for s in ("module calc_mod", "module linsolv_mod", "subroutine calcu()", "subroutine calcv()", "subroutine calcp()",
              "subroutine lisolv(istart, jstart, ni, nj, phi)"):
        assert s in raw, s
    # general family: external subroutines + tdma (buoyancy_cav.f90:122-433, 471-505)
    for s in ("subroutine calct()", "subroutine tdma(istart, jstart, ni, nj, phi)"):
        assert s in raw, s
    # heat programs: the steady 2-D driver routine with the reference's dummy list (fdm2d_steady_heat_cond.f90:68-69)
    assert "subroutine fdm2d_steady_heat_cond(xmin, xmax, ymin, ymax, tleft, tright, ttop, tbottom, t, nx, ny)" in raw
    # scalar-transport demos: the iterative line solvers (fvm2d_stag_point.f90:82,122; fvm2d_oblique_conv.f90:284)
    for s in ("subroutine tdma_we(aw, ae, as, an, ap, su, phi)", "subroutine tdma_sn(aw, ae, as, an, ap, su, phi)",
              "subroutine tdma_solver(aw, ae, as, an, ap, su, phi, ni, nj)"):
        assert s in raw, s
    # FD cavity programs: module cavity_mod with sipsol's dummy list (fdm_implicit_cavity.f90:29, fdm_explicit_cavity.f90:24)
    for s in ("module cavity_mod", "subroutine sipsol(i1, i2, j1, j2, aw, ae, as, an, ap, su, phi)"):
        assert s in raw, s
    # TEACH: module calc (calc.f90:6) + lisolv with leading dimensions (general.f90:294)
    for s in ("module calc ", "subroutine calcte()", "subroutine calced()",
              "subroutine lisolv(istart, jstart, ni, nj, it, jt, phi)"):
        assert s in raw, s


def test_integration_md_shows_the_same_types():
    """The interface block printed in INTEGRATION.md must not drift from the shim again."""
    shim_src = open(SHIM).read()
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    blocks = re.findall(r"```fortran\n(.*?)```", doc, flags=re.S)
    assert blocks
    ft_doc = {}
    for b in blocks:
        ft_doc.update(f_types(_strip_fortran(b)))
    ft = f_types(_strip_fortran(shim_src))
    assert "cfd_config" in ft_doc
    for name, comps in ft_doc.items():
        assert comps == ft[name], name
    for b in blocks:                                     # no positional constructors in the document either
        assert not re.search(r"cfd_config\(\s*\d", b)
