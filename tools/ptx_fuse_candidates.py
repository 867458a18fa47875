"""List PTX add/sub.f32 whose operand is a single-use mul.f32 (candidates for ptxas FMA fusion).
Dev tool used to work out which products the reference build contracts ('FMA contract')."""
import re, sys, collections
src = open(sys.argv[1]).read().splitlines()
ins = []
for l in src:
    l = l.strip()
    m = re.match(r'(@%p\d+\s+)?([a-z0-9_.]+)\s+(.*);', l)
    if m:
        ops = [o.strip() for o in m.group(3).split(',')]
        ins.append((m.group(2), ops, l))
uses = collections.Counter(); defs = {}
for op, ops, l in ins:
    if op.startswith('st.') or op.startswith('setp') :
        for o in ops: 
            for r in re.findall(r'%f\d+', o): uses[r] += 1
        continue
    if ops and re.fullmatch(r'%f\d+', ops[0]):
        defs.setdefault(ops[0], []).append((op, ops[1:], l))
        for o in ops[1:]:
            for r in re.findall(r'%f\d+', o): uses[r] += 1
    else:
        for o in ops:
            for r in re.findall(r'%f\d+', o): uses[r] += 1
for op, ops, l in ins:
    if op in ('add.f32', 'sub.f32'):
        for i, o in enumerate(ops[1:]):
            d = defs.get(o)
            if d and len(d) == 1 and d[0][0] == 'mul.f32':
                print(f"{l:50s} operand{i} {o} = {d[0][2]}  uses={uses[o]}")
