"""CPU: the jax.ffi face (ffi/spb_jax_ffi.cc, ffi/spb_jax.py) cannot be compiled or imported here (no jax, no xla/ffi
headers), so it is checked as SOURCE against include/spb200.h: every spb_* call has the arity the header declares, every
struct member it touches exists, every handler's Bind() chain matches its implementation's parameter list, and the
Python side registers exactly the symbols the C++ side defines and mirrors the ABI structs field for field."""
import ast
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CC = open(os.path.join(ROOT, "ffi", "spb_jax_ffi.cc")).read()
PY = open(os.path.join(ROOT, "ffi", "spb_jax.py")).read()
HDR = open(os.path.join(ROOT, "include", "spb200.h")).read()


def _strip_comments(src):
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.sub(r"//[^\n]*", "", src)


def _strip_strings(src):
    return re.sub(r'"(?:[^"\\]|\\.)*"', '""', src)


def _handlers(src):
    """[(symbol, impl, bind-chain text)] of every XLA_FFI_DEFINE_HANDLER_SYMBOL(...) in the C++ source"""
    out = []
    for m in re.finditer(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\(", src):
        body = _balanced(src, m.end() - 1)
        sym, impl, bind = (t.strip() for t in body.split(",", 2))
        out.append((sym, impl, bind))
    return out


def _split_args(s):
    out, depth, cur = [], 0, ""
    s = s.replace("->", ".")  # member access through a pointer is not a closing angle bracket
    for ch in s:
        if ch in "([{<":
            depth += 1
        elif ch in ")]}>":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def _balanced(src, start):
    """text between the parenthesis at src[start] and its match"""
    depth = 0
    for i in range(start, len(src)):
        if src[i] == "(":
            depth += 1
        elif src[i] == ")":
            depth -= 1
            if depth == 0:
                return src[start + 1:i]
    raise AssertionError("unbalanced")


def _header_protos():
    h = _strip_comments(HDR)
    protos = {}
    for m in re.finditer(r"\b(spb_\w+)\s*\(", h):
        name = m.group(1)
        args = _balanced(h, m.end() - 1)
        protos[name] = 0 if args.strip() in ("", "void") else len(_split_args(args))
    return protos


def _header_struct_fields():
    h = _strip_comments(HDR)
    out = {}
    for m in re.finditer(r"typedef struct (\w+) \{(.*?)\} \1;", h, flags=re.S):
        fields = set()
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for piece in decl.split(","):
                fields.add(re.sub(r"[\*\s]", " ", piece).split()[-1])
        out[m.group(1)] = fields
    return out


def test_every_c_abi_call_has_the_declared_arity():
    protos = _header_protos()
    cc = _strip_strings(_strip_comments(CC))
    calls = [(m.group(1), m.end() - 1) for m in re.finditer(r"\b(spb_[a-z_0-9]+)\s*\(", cc)]
    assert {"spb_block_fwd", "spb_block_bwd", "spb_block_ws_layout", "spb_norm_fwd", "spb_norm_bwd", "spb_rope_apply",
            "spb_last_error"} <= {c[0] for c in calls}
    for name, pos in calls:
        assert name in protos, f"{name} is not declared in include/spb200.h"
        args = _balanced(cc, pos)
        n = 0 if not args.strip() else len(_split_args(args))
        assert n == protos[name], f"{name}: {n} arguments in ffi/spb_jax_ffi.cc, {protos[name]} in the header"


def test_struct_members_exist_in_the_header():
    fields = _header_struct_fields()
    cc = _strip_comments(CC)
    for var, struct in (("d", "spb_block_desc"), ("p", "spb_block_params"), ("g", "spb_block_grads"), ("lay", "spb_ws_layout")):
        used = set(re.findall(rf"\b{var}(?:\.|->)(\w+)\b", cc))
        assert used, (var, struct)
        missing = used - fields[struct]
        assert not missing, f"{struct}: ffi source uses members the header does not declare: {missing}"
    # every member of the descriptor is initialised by make_desc (a forgotten one would be 0 by `= {}` - make it explicit)
    set_in_desc = set(re.findall(r"\bd\.(\w+)\s*=", cc))
    assert fields["spb_block_desc"] <= set_in_desc, fields["spb_block_desc"] - set_in_desc


def test_handlers_bind_what_their_implementations_take():
    cc = _strip_comments(CC)
    handlers = _handlers(cc)
    assert {h[0] for h in handlers} == {"SpbBlockFwd", "SpbBlockBwd", "SpbBlockDecode", "SpbRmsNorm", "SpbRmsNormBwd", "SpbRope"}
    for sym, impl, bind in handlers:
        n_bound = len(re.findall(r"\.(?:Ctx|Arg|Ret|Attr)<", bind))
        m = re.search(rf"ffi::Error {impl}\(", cc)
        assert m, impl
        n_params = len(_split_args(_balanced(cc, m.end() - 1)))
        assert n_bound == n_params, f"{sym}: Bind() lists {n_bound} items, {impl} takes {n_params} parameters"
        # order: context, args, results, attributes - and the implementation's types agree
        params = _split_args(_balanced(cc, m.end() - 1))
        kinds = re.findall(r"\.(Ctx|Arg|Ret|Attr)<", bind)
        for kind, prm in zip(kinds, params):
            if kind == "Ret":
                assert prm.startswith("ffi::Result<"), (sym, prm)
            elif kind == "Arg":
                assert prm.startswith("ffi::AnyBuffer") or prm.startswith("ffi::Buffer<"), (sym, prm)
            elif kind == "Ctx":
                assert prm.startswith("cudaStream_t"), (sym, prm)
    assert "cudaMalloc" not in cc and "cudaFree" not in cc and "cudaStreamSynchronize" not in cc and "cudaDeviceSynchronize" not in cc


def test_python_side_registers_the_defined_symbols_and_mirrors_the_abi_structs():
    tree = ast.parse(PY)  # the module is at least valid Python
    defined = set(re.findall(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\(\s*(\w+)", CC))
    registered = set(re.findall(r'\("spb_\w+",\s*"(\w+)"\)', PY))
    assert registered == defined, (registered, defined)
    targets = set(re.findall(r'\("(spb_\w+)",\s*"\w+"\)', PY))
    called = set(re.findall(r'ffi_call\(\s*"(\w+)"', PY))
    assert called == targets, (called, targets)
    names = {n.name for n in ast.walk(tree) if isinstance(n, ast.FunctionDef)}
    assert {"rms_norm", "apply_rotary_emb", "repeat_kv", "attention", "attention_block", "attention_decode", "_block_fwd", "_block_bwd",
            "register"} <= names
    # attributes passed by the Python side = attributes bound by the C++ side
    for target, sym in re.findall(r'\("(spb_\w+)",\s*"(\w+)"\)', PY):
        bind = [h[2] for h in _handlers(_strip_comments(CC)) if h[0] == sym][0]
        attrs = set(re.findall(r'\.Attr<[^>]+>\("(\w+)"\)', bind))
        for call in re.finditer(rf'ffi_call\(\s*"{target}"', PY):
            # the attribute keywords of this call site: from the ffi_call( to the end of the statement's second parenthesis group
            tail = PY[call.start():call.start() + 1600]
            first = _balanced(tail, tail.index("("))
            rest = tail[tail.index("(") + len(first) + 2:]
            second = _balanced(rest, rest.index("("))
            kw = set(re.findall(r"\b(\w+)=", second))
            assert kw == attrs, (target, kw, attrs)
    # struct mirrors: same field order as the ctypes binding the GPU tests use (which test_abi_symbols ties to the header)
    from solvingpapers_b200 import _lib as L
    ns = {}
    src = PY[PY.index("class _BlockDesc"):PY.index("def register")]
    exec("import ctypes\n" + src, ns)
    assert [f[0] for f in ns["_BlockDesc"]._fields_] == [f[0] for f in L.BlockDesc._fields_]
    assert [f[0] for f in ns["_WsLayout"]._fields_] == [f[0] for f in L.WsLayout._fields_]
    import ctypes
    assert ctypes.sizeof(ns["_BlockDesc"]) == ctypes.sizeof(L.BlockDesc) and ctypes.sizeof(ns["_WsLayout"]) == ctypes.sizeof(L.WsLayout)
    assert [f[0] for f in ns["_DecodeWs"]._fields_] == [f[0] for f in L.DecodeWs._fields_]


def test_shim_compiles_against_an_api_stand_in():
    """g++ -fsyntax-only against tests/ffi_standin (declarations of the public xla::ffi C++ API the shim uses - no XLA struct
    layouts) and the real include/spb200.h + CUDA runtime headers: catches type errors in the C-ABI calls."""
    import shutil
    import subprocess
    if shutil.which("g++") is None or not os.path.exists("/usr/local/cuda/include/cuda_runtime_api.h"):
        import pytest
        pytest.skip("needs g++ and the CUDA headers")
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-I" + os.path.join(ROOT, "tests", "ffi_standin"),
                        "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include", os.path.join(ROOT, "ffi", "spb_jax_ffi.cc")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
