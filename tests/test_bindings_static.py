"""CPU: the cgo shim under go/gpubackend/ is not compiled here (no Go toolchain), so at least keep it honest against the
header it binds: every `C.drx_*` call must name a function include/drynx_gpu.h declares and pass as many arguments as
the declaration has parameters. (Types are cgo's business; arity and names are what drifts when the ABI changes.)"""
import glob
import os
import re

ROOT = os.path.join(os.path.dirname(__file__), '..')


def _split_top_level(s):
    out, depth, cur = [], 0, ''
    for ch in s:
        if ch in '([{':
            depth += 1
        elif ch in ')]}':
            depth -= 1
        if ch == ',' and depth == 0:
            out.append(cur)
            cur = ''
        else:
            cur += ch
    if cur.strip():
        out.append(cur)
    return out


def _header_arity():
    src = open(os.path.join(ROOT, 'include', 'drynx_gpu.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    src = re.sub(r'//.*', '', src)
    decl = {}
    for m in re.finditer(r'\b(drx_\w+)\s*\(([^;{]*?)\)\s*;', src, flags=re.S):
        params = m.group(2).strip()
        decl[m.group(1)] = 0 if params in ('', 'void') else len(_split_top_level(params))
    return decl


def _go_calls():
    calls = []
    for path in sorted(glob.glob(os.path.join(ROOT, 'go', 'gpubackend', '*.go'))) + [os.path.join(ROOT, 'INTEGRATION.md')]:
        src = open(path).read()
        src = re.sub(r'//.*', '', src)
        for m in re.finditer(r'\bC\.(drx_\w+)\s*\(', src):
            depth, i = 1, m.end()
            while depth:
                depth += {'(': 1, ')': -1}.get(src[i], 0)
                i += 1
            args = src[m.end():i - 1].strip()
            calls.append((os.path.basename(path), m.group(1), len(_split_top_level(args)) if args else 0))
    return calls


def test_go_shim_calls_match_the_header():
    decl = _header_arity()
    calls = _go_calls()
    assert len(decl) > 40 and len(calls) > 25
    for path, name, nargs in calls:
        assert name in decl, '%s calls %s, which include/drynx_gpu.h does not declare' % (path, name)
        assert nargs == decl[name], '%s: %s called with %d arguments, declared with %d' % (path, name, nargs, decl[name])


def test_go_shim_covers_the_hot_path_entry_points():
    used = {name for _, name, _ in _go_calls()}
    for name in ('drx_init', 'drx_free', 'drx_set_pubkey', 'drx_encrypt', 'drx_aggregate', 'drx_dlog_table', 'drx_decrypt',
                 'drx_rp_make_sigs', 'drx_rp_set_sigs', 'drx_rp_create', 'drx_rp_verify', 'drx_keyswitch',
                 'drx_ct_scalar_mul', 'drx_shuffle_rerandomize', 'drx_schnorr_verify', 'drx_rep_commit',
                 'drx_rep_respond', 'drx_rep_verify'):
        assert name in used, name


def test_ctypes_binding_has_the_header_arity():
    """drynx_b200/_lib.py binds every declared function with as many argtypes as the declaration has parameters."""
    import sys
    sys.path.insert(0, ROOT)
    from drynx_b200 import _lib
    decl = _header_arity()
    for name, (_, argtypes) in _lib.SIGNATURES.items():
        assert name in decl, name
        assert len(argtypes) == decl[name], '%s: %d argtypes, %d parameters' % (name, len(argtypes), decl[name])
    assert set(decl) == set(_lib.SIGNATURES)
