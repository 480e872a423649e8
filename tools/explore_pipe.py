"""Compile-only sweep of sgemm_pipe_kernel variants: builds a cubin per batch and reports, per variant, registers, spills and the
static FFMA2 operand statistics (fresh register reads per FFMA2 = reads not served by the operand-reuse cache).  No GPU needed."""
import itertools, os, re, subprocess, sys, collections, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "guda_b200", "csrc")

def analyze(sass):
    res = {}
    for f in re.split(r'\n\s+Function : ', sass):
        name = f.split('\n', 1)[0]
        m = re.search(r'PipeCfgI(.*?)EEELi0ELb0', name)
        if not m: continue
        cfg = tuple(int(v) for v in re.sub(r'[LiEb]', ' ', m.group(1)).split())
        prev = None; c = collections.Counter()
        for line in f.split('\n'):
            mm = re.match(r'\s+/\*[0-9a-f]+\*/\s+(.*?);', line)
            if not mm: continue
            txt = mm.group(1)
            if not txt.startswith('FFMA2'):
                if not txt.startswith('LDS'): prev = None
                continue
            ops = [o.strip() for o in txt[len('FFMA2'):].split(',')]
            cur = {}; fresh = 0
            for slot, o in enumerate(ops[1:4]):
                r = int(re.match(r'R(\d+)', o).group(1)); wide = 'F32x2' in o
                if '.reuse' in o: cur[slot] = (r, wide)
                if not (prev is not None and prev.get(slot) == (r, wide)): fresh += 2 if wide else 1
            c['n'] += 1; c['fresh'] += fresh; prev = cur
        res[cfg] = c['fresh'] / max(c['n'], 1), c['n']
    return res

def run(variants):
    src = '#include "sgemm_pipe.cuh"\nnamespace guda {\n'
    for v in variants:
        src += 'template __global__ void sgemm_pipe_kernel<PipeCfg<%s>, EPI_NONE, false>(const __grid_constant__ CUtensorMap, const __grid_constant__ CUtensorMap, GemmParams<float>, const unsigned char*, int, int, int, ScatterParams<float, false>);\n' % ', '.join(str(x).lower() for x in v)
    src += '}\n'
    with tempfile.TemporaryDirectory() as d:
        cu = os.path.join(d, 'x.cu'); cubin = os.path.join(d, 'x.cubin')
        open(cu, 'w').write(src)
        p = subprocess.run(['nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-std=c++17', '-fmad=false', '-Xptxas', '-v', '-I', CSRC, '-cubin', '-o', cubin, cu], capture_output=True, text=True)
        if p.returncode: print(p.stderr[-2000:]); sys.exit(1)
        info = {}
        blocks = re.split(r"ptxas info\s+: Compiling entry function '", p.stderr)
        for b in blocks[1:]:
            m = re.search(r'PipeCfgI(.*?)EEELi0ELb0', b)
            cfg = tuple(int(v) for v in re.sub(r'[LiEb]', ' ', m.group(1)).split())
            regs = int(re.search(r'Used (\d+) registers', b).group(1))
            sp = re.search(r'(\d+) bytes spill stores', b)
            info[cfg] = (regs, int(sp.group(1)) if sp else 0)
        sass = subprocess.run(['cuobjdump', '-sass', cubin], capture_output=True, text=True).stdout
        an = analyze(sass)
    return info, an

if __name__ == '__main__':
    # TM TN LY LX WM WN BK ST MINB ORDER DRY PAIRSEL
    shapes = [(16, 8, 2, 16, 4, 1, 16, 4, 2), (8, 16, 2, 16, 4, 1, 16, 4, 2), (8, 8, 2, 16, 4, 1, 32, 3, 2), (8, 8, 2, 16, 4, 1, 16, 4, 2), (16, 8, 1, 32, 4, 1, 16, 4, 2)]
    if len(sys.argv) > 1:
        shapes = [tuple(int(x) for x in a.split(',')) for a in sys.argv[1:]]
    variants = [s + (o, 0, ps, a) for s in shapes for o in (1, 2) for ps in (0, 1) for a in (0, 1)]
    info, an = run(variants)
    for v in variants:
        key = tuple(int(x) for x in v)
        f, n = an.get(key, (0, 0)); regs, spill = info.get(key, (0, 0))
        print('%-44s regs %3d spill %3d  FFMA2 %4d  fresh reads/FFMA2 %.2f' % (' '.join(map(str, v)), regs, spill, n, f))
