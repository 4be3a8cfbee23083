"""profiles/sass_k_fast_<tag>.md: the column loops of the trace kernels as ptxas built them (cuobjdump -sass of the library):
   python scripts/sass_excerpt.py <tag>"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, 'custom_porechop_abi_b200', 'cpp_functions.so')
tag = sys.argv[1]
sass = subprocess.run(['cuobjdump', '-sass', SO], capture_output=True, text=True).stdout
funcs = re.split(r'\n\s*Function : ', sass)


def loops(body):
    ins = [(int(m.group(1), 16), m.group(2)) for m in (re.match(r'\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);', l) for l in body.split('\n')) if m]
    idx = {a: i for i, (a, _) in enumerate(ins)}
    out = []
    for i, (a, t) in enumerate(ins):
        m = re.search(r'\bBRA\b.*?(0x[0-9a-f]+)', t)
        if m and int(m.group(1), 16) <= a and int(m.group(1), 16) in idx:
            out.append(ins[idx[int(m.group(1), 16)]:i + 1])
    return out


def hist(body):
    h = collections.Counter(re.sub(r'^@!?U?P\d+\s+', '', t).split()[0] for _, t in body)
    return ', '.join('%s %d' % kv for kv in h.most_common(16))


def pick(name_part, key, max_len):
    for f in funcs[1:]:
        name = f.split('\n', 1)[0]
        if name_part in name:
            cands = [b for b in loops(f) if len(b) <= max_len]
            best = max(cands, key=lambda b: sum(key in t for _, t in b))
            return name, best
    raise SystemExit('kernel not found: ' + name_part)


lines = ['# SASS of the trace kernels\' column loops (`cuobjdump -sass custom_porechop_abi_b200/cpp_functions.so`, build `%s`)' % tag, '']
for title, part, key, max_len, rows, cols in (
        ('pass 1, 24-base exact class: the lean column loop (2 columns x 24 rows per trip)', 'k_fastILi1ELi24ELb1E', 'VIMNMX3', 400, 24, 2),
        ('pass 2, 24-base exact class: one flag column (24 rows per trip)', 'k_roundsILi24ELb1E', 'PRMT', 600, 24, 1),
        ('pass 2: the walk, both tracebacks of a thread per trip', 'k_roundsILi24ELb1E', 'VIMNMX.S16x2', 120, 0, 0)):
    name, body = pick(part, key, max_len)
    lines += ['## ' + title, '', '`%s`, loop 0x%x..0x%x: %d instructions%s' % (
        name.strip(), body[0][0], body[-1][0], len(body), (' = %.2f per cell pair' % (len(body) / (rows * cols))) if rows else ''), '',
        'opcode mix: ' + hist(body), '', '```']
    lines += ['%04x  %s' % (a, t.strip()) for a, t in body[:(72 if rows else 100)]]
    if len(body) > 72 and rows:
        lines.append('...  (%d more of the same pattern)' % (len(body) - 72))
    lines += ['```', '']
lines += ['Per row of the lean loop: `LDS` (substitution scores of the two tasks from the warp\'s table), `IMAD.IADD` (diagonal + score, on the',
          'FMA pipe), `VIADDMNMX.S16x2` (H), `VIADDMNMX.S16x2` (V), `VIMNMX3.S16x2` (M) -- the three DPX instructions are the ALU-pipe work',
          'the integer roofline counts.  The flag column adds, per row, four `IADD3` (a + 0x8000 - b: bit 15 = a >= b), two `PRMT` in',
          'sign-replicate mode and two `LOP3` that pack the eight flag bits, and `VIMNMX.S16x2` twice instead of `VIMNMX3`.', '']
open(os.path.join(ROOT, 'profiles', 'sass_k_fast_%s.md' % tag), 'w').write('\n'.join(lines))
print('wrote profiles/sass_k_fast_%s.md' % tag)
