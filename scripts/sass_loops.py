"""Opcode histogram of the loops of one kernel in a cuobjdump -sass listing: python scripts/sass_loops.py <file.sass> [min_len]
A loop = a backward branch; prints, for every loop body, its length and opcode mix."""
import re, sys, collections
lines = open(sys.argv[1]).read().split('\n')
minlen = int(sys.argv[2]) if len(sys.argv) > 2 else 40
ins = []
for l in lines:
    m = re.match(r'\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);', l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2)))
addr_index = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    m = re.search(r'\bBRA\b.*?(0x[0-9a-f]+)', t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt <= a and tgt in addr_index and i - addr_index[tgt] >= minlen:
            body = ins[addr_index[tgt]:i + 1]
            h = collections.Counter()
            for _, x in body:
                x = re.sub(r'^@!?U?P\d+\s+', '', x)
                op = x.split()[0]
                h[op] += 1
            print('loop 0x%x..0x%x  %d instr' % (tgt, a, len(body)))
            print('   ' + ', '.join('%s %d' % kv for kv in h.most_common(24)))
