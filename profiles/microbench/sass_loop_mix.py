"""Count SASS instructions in the largest loop of each kernel of a cuobjdump -sass dump (stdin or file)."""
import re, collections, sys
txt = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
funcs = re.split(r'\n\s*Function : ', txt)[1:]
for f in funcs:
    name = f.split('\n')[0]
    ins = [(int(m.group(1), 16), m.group(2)) for m in re.finditer(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);', f)]
    best = None
    for a, t in ins:
        m = re.search(r'BRA\S*\s+(?:.*?)0x([0-9a-f]+)', t)
        if m and int(m.group(1), 16) < a:
            tgt = int(m.group(1), 16)
            if best is None or a - tgt > best[1] - best[0]:
                best = (tgt, a)
    body = [t for a, t in ins if best is None or best[0] <= a <= best[1]]
    c = collections.Counter()
    for t in body:
        t = re.sub(r'^@!?U?P\d+\s+', '', t)
        c[t.split()[0]] += 1
    print(name, len(body), dict(c.most_common(14)))
