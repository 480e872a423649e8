import re, subprocess, sys, collections
lib = 'guda_b200/lib/libguda_b200.so'
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
funcs = re.split(r'\n\s+Function : ', out)
res = {}
for f in funcs:
    name = f.split('\n', 1)[0]
    if 'sgemm_pipe_kernel' not in name or 'Li0ELb0EEEv' not in name: continue
    m = re.search(r'PipeCfgI(.*?)EEELi0ELb0', name)
    cfg = ' '.join(re.sub(r'[LiEb]', ' ', m.group(1)).split())
    prev = None
    c = collections.Counter()
    for line in f.split('\n'):
        mm = re.match(r'\s+/\*[0-9a-f]+\*/\s+(.*?);', line)
        if not mm: continue
        txt = mm.group(1)
        if not txt.startswith('FFMA2'):
            prev = None if not txt.startswith('LDS') else prev   # model: any non-LDS instruction clears the reuse cache
            if txt.startswith('LDS'): c['lds'] += 1
            continue
        ops = [o.strip() for o in txt[len('FFMA2'):].split(',')]
        dst = int(re.match(r'R(\d+)', ops[0]).group(1))
        cur = {}; fresh = []; nre = 0
        for slot, o in enumerate(ops[1:4]):
            reg = re.match(r'R(\d+)', o)
            r = int(reg.group(1)); wide = 'F32x2' in o
            cached = prev is not None and prev.get(slot) == (r, wide)
            if '.reuse' in o: cur[slot] = (r, wide)
            if cached: nre += 1
            else: fresh += [r, r + 1] if wide else [r]
        c['n'] += 1
        ev = sum(1 for r in fresh if r % 2 == 0); od = len(fresh) - ev
        if max(ev, od) > 2: c['m2bank'] += 1
        b4 = collections.Counter(r % 4 for r in fresh)
        if max(b4.values()) > 1: c['m4bank_any'] += 1
        c['m4bank_extra'] += sum(v - 1 for v in b4.values())
        if nre == 0: c['noreuse'] += 1
        src3 = int(re.match(r'R(\d+)', ops[3]).group(1))
        if dst != src3: c['renamed'] += 1
        c['fresh_regs'] += len(fresh)
        prev = cur
    res[cfg] = c
for cfg, c in sorted(res.items()):
    n = c['n']
    print('%-36s n=%d 2bank>2: %4.1f%%  4bank-any: %4.1f%%  4bank-extra/instr: %.2f  noreuse: %4.1f%%  renamed: %4.1f%%  fresh regs/instr %.2f' % (
        cfg, n, 100*c['m2bank']/n, 100*c['m4bank_any']/n, c['m4bank_extra']/n, 100*c['noreuse']/n, 100*c['renamed']/n, c['fresh_regs']/n))
