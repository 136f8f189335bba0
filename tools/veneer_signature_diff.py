#!/usr/bin/env python
"""Compares the parameter lists of every `*_kernel` / `*_task_gpu` / `*_wrapper` function declared under the
reference's lib/kernels/include/kernels with the C++ veneer (flexflow-train_b200/include/ffk_cpp), namespace by
namespace, modulo const placement and parameter names. Prints MISSING / DIFF lines and returns their count.
Needs the reference checkout; used by tests/test_abi_cpu.py.

    python tools/veneer_signature_diff.py [/root/reference]
"""
import re, sys
from pathlib import Path

def strip(t):
    t = re.sub(r'//[^\n]*', ' ', t); t = re.sub(r'/\*.*?\*/', ' ', t, flags=re.S)
    t = re.sub(r'^\s*#.*$', ' ', t, flags=re.M)
    return t

def split_params(p):
    out, depth, cur = [], 0, ''
    for ch in p:
        if ch in '<(': depth += 1
        if ch in '>)': depth -= 1
        if ch == ',' and depth == 0:
            out.append(cur); cur = ''
        else: cur += ch
    if cur.strip(): out.append(cur)
    return out

def norm_type(param):
    param = param.split('=')[0]
    param = ' '.join(param.split())
    # drop the parameter name: last identifier if preceded by type chars
    m = re.match(r'^(.*?[\s\*&>])([A-Za-z_][A-Za-z0-9_]*)$', param)
    if m and m.group(1).strip() and m.group(1).strip() not in ('const', 'unsigned'):
        t = m.group(1)
    else:
        t = param
    t = t.replace('std::', '').replace('FlexFlow::', '')
    t = re.sub(r'\s+', '', t)
    t = t.replace('const', '')   # const placement differs; compare modulo const
    return t

def functions(text):
    """yield (namespace path, name, [param types]) for declarations/definitions"""
    text = strip(text)
    # track namespaces with a simple stack
    tokens = re.finditer(r'namespace\s+([A-Za-z_][A-Za-z0-9_:]*)\s*\{|\{|\}|([A-Za-z_][A-Za-z0-9_]*)\s*\(', text)
    stack = []
    res = []
    i = 0
    pos = 0
    for m in re.finditer(r'namespace\s+([A-Za-z_][A-Za-z0-9_:]*)\s*\{|(\{)|(\})', text):
        pass
    # simpler: iterate chars
    ns_stack = []
    brace_stack = []
    idx = 0
    n = len(text)
    pat_ns = re.compile(r'namespace\s+([A-Za-z_][A-Za-z0-9_:]*)\s*\{')
    pat_fn = re.compile(r'\b([a-z_][A-Za-z0-9_]*)\s*\(')
    while idx < n:
        m = pat_ns.match(text, idx)
        if m:
            brace_stack.append(('ns', m.group(1))); idx = m.end(); continue
        ch = text[idx]
        if ch == '{':
            brace_stack.append(('blk', None)); idx += 1; continue
        if ch == '}':
            if brace_stack: brace_stack.pop()
            idx += 1; continue
        m = pat_fn.match(text, idx)
        if m and all(k == 'ns' for k, _ in brace_stack):
            name = m.group(1)
            # find matching paren
            j = m.end(); depth = 1
            while j < n and depth:
                if text[j] == '(': depth += 1
                elif text[j] == ')': depth -= 1
                j += 1
            params = text[m.end():j-1]
            ns = '::'.join(v for k, v in brace_stack if k == 'ns')
            if name.endswith('_kernel') or name.endswith('_task_gpu') or name.endswith('_wrapper'):
                res.append((ns.replace('FlexFlow::', '').replace('FlexFlow', ''), name, [norm_type(p) for p in split_params(params)]))
            idx = j; continue
        idx += 1
    return res

def compare(ref_root='/root/reference', out=print):
    ref = Path(ref_root) / 'lib' / 'kernels' / 'include' / 'kernels'
    ven = Path(__file__).resolve().parents[1] / 'flexflow-train_b200' / 'include' / 'ffk_cpp'
    return _compare(ref, ven, out)


def _compare(ref, ven, out):
  R = {}
  for h in sorted(ref.glob('*.h')):
      for ns, name, params in functions(h.read_text()):
          R.setdefault((ns, name), []).append((h.name, params))
  V = {}
  for h in sorted(ven.glob('*.h')):
      for ns, name, params in functions(h.read_text()):
          V.setdefault((ns, name), []).append((h.name, params))
  bad = 0
  for key, decls in sorted(R.items()):
      if key not in V:
          out(f'MISSING {key} {decls[0][0]}'); bad += 1; continue
      vs = [p for _, p in V[key]]
      for hname, params in decls:
          if params not in vs:
              out(f'DIFF {key} {hname}'); out(f'   ref: {params}')
              for v in vs: out(f'   ven: {v}')
              bad += 1
  out(f'checked {len(R)} bad {bad}')
  return len(R), bad


def compare_call_sites(ref_root='/root/reference', out=print):
    """The callers' view: every `profile(<kernel>, settings, "fmt", args...)` in
    lib/local-execution/src/ops/*.cc must find <kernel> in the veneer — in one of the namespaces the
    file pulls in with `using namespace FlexFlow::Kernels::<Op>`, or in FlexFlow — with exactly
    len(args) + 1 parameters (profile() prepends the stream). Returns (call sites, mismatches)."""
    ven = Path(__file__).resolve().parents[1] / 'flexflow-train_b200' / 'include' / 'ffk_cpp'
    V = {}
    for h in sorted(ven.glob('*.h')):
        for ns, name, params in functions(h.read_text()):
            V.setdefault((ns, name), []).append(params)
    total = bad = 0
    for f in sorted((Path(ref_root) / 'lib' / 'local-execution' / 'src' / 'ops').glob('*.cc')):
        t = re.sub(r'//[^\n]*', ' ', f.read_text())
        using = re.findall(r'using namespace FlexFlow::Kernels::([A-Za-z0-9_]+)', t)
        for m in re.finditer(r'\bprofile\s*\(', t):
            i, depth = m.end(), 1
            while depth and i < len(t):
                depth += {'(': 1, ')': -1}.get(t[i], 0)
                i += 1
            args = split_params(t[m.end():i - 1])
            fn, nargs = args[0].strip(), len(args) - 3
            total += 1
            cands = [c for ns in using for c in V.get((f'Kernels::{ns}', fn), [])] + V.get(('', fn), [])
            if not cands:
                out(f'{f.name}: {fn} not found in the veneer (namespaces {using})')
                bad += 1
            elif not any(len(c) - 1 == nargs for c in cands):
                out(f'{f.name}: {fn} is called with {nargs} arguments, the veneer takes {[len(c) - 1 for c in cands]}')
                bad += 1
        # init_kernel is called directly (no stream, no profile())
        for m in re.finditer(r'(?<![A-Za-z_])init_kernel\s*\(', t):
            i, depth = m.end(), 1
            while depth and i < len(t):
                depth += {'(': 1, ')': -1}.get(t[i], 0)
                i += 1
            nargs = len(split_params(t[m.end():i - 1]))
            total += 1
            cands = [c for ns in using for c in V.get((f'Kernels::{ns}', 'init_kernel'), [])]
            if not any(len(c) == nargs for c in cands):
                out(f'{f.name}: init_kernel is called with {nargs} arguments, the veneer takes {[len(c) for c in cands]}')
                bad += 1
    out(f'call sites {total} mismatches {bad}')
    return total, bad


if __name__ == '__main__':
    n, bad = compare(*(sys.argv[1:2]))
    n2, bad2 = compare_call_sites(*(sys.argv[1:2]))
    sys.exit(1 if bad or bad2 else 0)
