"""Static check of the FFMA2 stream: how many FFMA2 read 3 registers from one even/odd bank (no operand in the reuse cache)."""
import re, subprocess, sys
lib = 'guda_b200/lib/libguda_b200.so'
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
funcs = re.split(r'\n\s+Function : ', out)
for f in funcs:
    name = f.split('\n', 1)[0]
    if 'sgemm_pipe_kernel' not in name or 'Li0ELb0EEEv' not in name: continue
    m = re.search(r'PipeCfgI(.*?)EEELi0ELb0', name)
    cfg = re.sub(r'[LiEb]', ' ', m.group(1)).split() if m else name
    prev = None; total = conflict = reused = 0
    for line in f.split('\n'):
        mm = re.match(r'\s+/\*[0-9a-f]+\*/\s+(.*?);', line)
        if not mm: continue
        txt = mm.group(1)
        if not txt.startswith('FFMA2'):
            if not txt.startswith('LDS') and not txt.startswith('NOP'): prev = None if txt.split()[0] in ('BRA','BSYNC','BSSY','SYNCS','WARPSYNC','BAR') else prev
            continue
        ops = [o.strip() for o in txt[len('FFMA2'):].split(',')]
        srcs = ops[1:4]
        cur = {}
        fresh = []
        for slot, o in enumerate(srcs):
            reg = re.match(r'R(\d+)', o)
            if not reg: continue
            r = int(reg.group(1)); wide = 'F32x2' in o
            cached = prev is not None and prev.get(slot) == (r, wide)
            if '.reuse' in o: cur[slot] = (r, wide)
            if cached: reused += 1
            else: fresh += [r, r + 1] if wide else [r]
        ev = sum(1 for r in fresh if r % 2 == 0); od = len(fresh) - ev
        total += 1
        if ev > 2 or od > 2: conflict += 1
        prev = cur
    print('%-40s FFMA2 %5d  bank-conflicting %5d (%.1f%%)  reuse hits %d' % (' '.join(cfg), total, conflict, 100.0 * conflict / max(total, 1), reused))
