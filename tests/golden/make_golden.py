#!/usr/bin/env python
"""Extract the reference's own golden vectors into JSON fixtures.

Run in the build container (needs /root/reference, which does NOT exist on the
GPU box):   python tests/golden/make_golden.py
Outputs (committed): tests/golden/{dgemm_cases,level1_double,baseline_kat,gelu_kat}.json

Sources (reference file:line):
  blas/testblas/dgemm.go:26-215          DgemmCases (7 cases; each run NN/TN/NT/TT)
  blas/testblas/level1double.go:33-...   DoubleOneVectorCases (Dscal answers)
  blas/testblas/level1double.go:552-...  DoubleTwoVectorCases (Daxpy answers, Ddot)
  baseline_20250807_221713.json          recorded AXPY_1000 / DOT_1000 / GEMM_128 results
  fused_gelu_test.go:10-22               GELU point values
This is a data extraction: a small parser for Go composite literals, no arithmetic.
"""
import json
import os
import re
import sys

REF = os.environ.get("GUDA_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))

TOKEN = re.compile(r"""
    (?P<ws>\s+|//[^\n]*)            |
    (?P<str>"(?:[^"\\]|\\.)*")      |
    (?P<num>[-+]?(?:\d+\.?\d*(?:[eE][-+]?\d+)?|\.\d+(?:[eE][-+]?\d+)?)) |
    (?P<id>[A-Za-z_][A-Za-z_0-9.]*) |
    (?P<type>\[\d*\](?:\[\d*\])*[A-Za-z_][A-Za-z_0-9.]*) |
    (?P<p>[{}(),:*/+-])
""", re.X)


def tokenize(src):
    pos, out = 0, []
    while pos < len(src):
        m = TOKEN.match(src, pos)
        if not m:
            raise ValueError("bad token at %r" % src[pos:pos + 40])
        pos = m.end()
        if m.lastgroup != "ws":
            out.append((m.lastgroup, m.group()))
    return out


class Parser:
    def __init__(self, toks):
        self.t, self.i = toks, 0

    def peek(self, k=0):
        return self.t[self.i + k] if self.i + k < len(self.t) else (None, None)

    def eat(self):
        tok = self.t[self.i]; self.i += 1
        return tok

    def value(self):
        kind, text = self.peek()
        if kind == "type":              # []float64{...}
            self.eat()
            return self.value()
        if kind == "p" and text == "{":
            return self.composite()
        if kind == "str":
            self.eat()
            return json.loads(text)
        # an expression up to the next top-level , or }
        return self.expr()

    def expr(self):
        depth, parts = 0, []
        while True:
            kind, text = self.peek()
            if kind is None:
                break
            if depth == 0 and kind == "p" and text in ",}":
                break
            if kind == "p" and text == "(":
                depth += 1
            if kind == "p" and text == ")":
                depth -= 1
            if kind == "id" and self.peek(1) == ("p", "{"):      # TypeName{...}
                self.eat()
                return self.composite()
            parts.append(text); self.eat()
        s = " ".join(parts)
        if s in ("true", "false"):
            return s == "true"
        try:
            return int(s)
        except ValueError:
            pass
        try:
            return float(s)
        except ValueError:
            pass
        compact = s.replace(" ", "")
        special = {"math.Inf(1)": float("inf"), "math.Inf(-1)": float("-inf"), "math.NaN()": float("nan"),
                   "-math.Inf(1)": float("-inf")}
        if compact in special:
            return special[compact]
        return {"_expr": s}

    def composite(self):
        assert self.eat() == ("p", "{")
        keyed = self.peek()[0] == "id" and self.peek(1) == ("p", ":")
        res = {} if keyed else []
        while self.peek() != ("p", "}"):
            if keyed:
                key = self.eat()[1]; self.eat()
                res[key] = self.value()
            else:
                res.append(self.value())
            if self.peek() == ("p", ","):
                self.eat()
        self.eat()
        return res


def extract_var(path, varname):
    src = open(path).read()
    m = re.search(r"var\s+%s\s*=\s*\[\][A-Za-z]+\s*\{" % varname, src)
    start = m.end() - 1
    depth, j = 0, start
    while True:
        ch = src[j]
        if ch == '"':
            j = src.index('"', j + 1)
        elif ch == "{":
            depth += 1
        elif ch == "}":
            depth -= 1
            if depth == 0:
                break
        j += 1
    return Parser(tokenize(src[start:j + 1])).composite()


def extract_struct_slice(path, varname):
    """`var name = []struct { ...fields... }{ ...elements... }`: skip the anonymous struct type, parse the element list."""
    src = open(path).read()
    m = re.search(r"var\s+%s\s*=\s*\[\]struct\s*\{" % varname, src)
    j = src.index("}", m.end())                       # flat field list: the first closing brace ends the type
    start = src.index("{", j)
    depth, j = 0, start
    while True:
        ch = src[j]
        if ch == "{":
            depth += 1
        elif ch == "}":
            depth -= 1
            if depth == 0:
                break
        j += 1
    body = re.sub(r"//[^\n]*", "", src[start:j + 1])
    return Parser(tokenize(body)).composite()


def Fraction_eval(expr):
    """Exact rational evaluation of a Go constant expression made of decimal literals, + - * and parentheses."""
    from fractions import Fraction
    expr = re.sub(r"(\d+\.?\d*)", lambda m: "Fraction('%s')" % m.group(1), expr)
    return eval(expr, {"Fraction": Fraction})


def extract_reduce_cases(path):
    """reduce_test.go: testSum/testMax/testMin/testArgMax/testArgMin/testSumSquares each hold a table of
    {"name", []float32{...} | makeSequence(k), expected} rows."""
    src = open(path).read()
    out = {}
    for fn in ("testSum", "testMax", "testMin", "testArgMax", "testArgMin", "testSumSquares"):
        body = src[src.index("func %s(" % fn):]
        body = body[:body.index("\nfunc ", 10)]
        rows = []
        for name, inp, exp in re.findall(r'\{"(\w+)",\s*(\[\]float32\{[^}]*\}|makeSequence\(\d+\)),\s*([^}]+?)\},?\s*(?://[^\n]*)?\n', body):
            if inp.startswith("makeSequence"):
                vals = [float(i) for i in range(int(re.search(r"\d+", inp).group()))]
            else:
                vals = [float(v) for v in re.findall(r"-?\d+\.?\d*", inp[inp.index("{"):])]
            exp = exp.strip()
            if "math.Inf(-1)" in exp: e = "-inf"
            elif "math.Inf(1)" in exp: e = "inf"
            else: e = float(exp)
            rows.append({"name": name, "input": vals, "expected": e})
        out[fn] = rows
    return out


def extract_conv_cases(path):
    """conv_test.go: each TestConv2D* holds a ConvParams literal, input/kernel[/bias] float32 literals and `expected`."""
    src = open(path).read()
    src = src[:src.index("func BenchmarkConv2D")]
    cases = []
    for name, body in re.findall(r"func (TestConv2D\w+)\(t \*testing\.T\) \{(.*?)\n\}\n", src, flags=re.S):
        body = re.sub(r"//[^\n]*", "", body)
        params = {k: (v == "true" if v in ("true", "false") else int(v))
                  for k, v in re.findall(r"(\w+):\s*(\d+|true|false),", body[body.index("&ConvParams{"):body.index("}", body.index("&ConvParams{"))])}
        def floats(var):
            m = re.search(var + r"\s*:=\s*\[\]float32\{(.*?)\}", body, flags=re.S)
            return [float(x) for x in re.findall(r"-?\d+\.?\d*", m.group(1))] if m else None
        exp = floats("expected")
        if exp is None:
            exp = [float(re.search(r"expected\s*:=\s*float32\((-?\d+\.?\d*)\)", body).group(1))]
        cases.append({"name": name, "params": params, "input": floats("input"), "kernel": floats("kernel"), "bias": floats("bias"),
                      "expected": exp})
    return cases


def main():
    dg = extract_var(os.path.join(REF, "blas/testblas/dgemm.go"), "DgemmCases")
    assert len(dg) == 7, len(dg)
    json.dump({"source": "blas/testblas/dgemm.go:26-215", "cases": dg},
              open(os.path.join(OUT, "dgemm_cases.json"), "w"), indent=1)

    one = extract_var(os.path.join(REF, "blas/testblas/level1double.go"), "DoubleOneVectorCases")
    two = extract_var(os.path.join(REF, "blas/testblas/level1double.go"), "DoubleTwoVectorCases")
    keep1 = [{k: c[k] for k in ("Name", "X", "Incx", "N", "Panic", "DscalCases") if k in c} for c in one]
    keep2 = [{k: c[k] for k in ("Name", "X", "Y", "Incx", "Incy", "N", "Panic", "DaxpyCases", "DdotAns") if k in c}
             for c in two]
    json.dump({"source": "blas/testblas/level1double.go:33-,552-", "one_vector": keep1, "two_vector": keep2},
              open(os.path.join(OUT, "level1_double.json"), "w"), indent=1)

    base = json.load(open(os.path.join(REF, "baseline_20250807_221713.json")))
    kat = {b["test_name"]: {"checksum": b["checksum"], "sample_values": b["sample_values"]} for b in base}
    json.dump({"source": "baseline_20250807_221713.json (generator baseline_test.go:34-152)", "results": kat},
              open(os.path.join(OUT, "baseline_kat.json"), "w"), indent=1)

    src = open(os.path.join(REF, "fused_gelu_test.go")).read()
    pts = re.findall(r"\{\s*(-?\d+\.?\d*)\s*,\s*(-?\d+\.?\d*)\s*,\s*([-\d.e]+)\s*\}", src[:src.index("func TestAddBiasGELU")])
    json.dump({"source": "fused_gelu_test.go:10-22 (input, expected, tol)", "points": [[float(a), float(b), float(c)] for a, b, c in pts]},
              open(os.path.join(OUT, "gelu_kat.json"), "w"), indent=1)
    conv = extract_conv_cases(os.path.join(REF, "conv_test.go"))
    assert len(conv) == 4, len(conv)
    json.dump({"source": "conv_test.go:8-260 (TestConv2DBasic/WithBias/Stride/MultiChannel: params, inputs, expected, tol 1e-5)",
               "cases": conv}, open(os.path.join(OUT, "conv_cases.json"), "w"), indent=1)
    gemv = extract_var(os.path.join(REF, "compute/asm/f32/gemv_test.go"), "SgemvCases")
    ger = extract_struct_slice(os.path.join(REF, "compute/asm/f32/ge_test.go"), "gerTests")
    assert len(gemv) == 7 and len(ger) >= 8, (len(gemv), len(ger))
    json.dump({"source": "compute/asm/f32/gemv_test.go:35-506 (SgemvCases; harness :508-615, tol 1e-6 abs-or-rel); "
                         "compute/asm/f32/ge_test.go:14-144 (gerTests; harness :146-219, alpha 1 unit / 3.5 strided)",
               "gemv": gemv, "ger": ger}, open(os.path.join(OUT, "level2_cases.json"), "w"), indent=1)
    dgv = extract_var(os.path.join(REF, "blas/testblas/dgemv.go"), "DgemvCases")
    for c in dgv:
        c["tA"] = c["tA"]["_expr"] if isinstance(c["tA"], dict) else c["tA"]
        for sub in c["Subcases"]:            # a few answers are written as constant expressions ("40.8 * 8 - 6 * 7"): Go folds them exactly
            sub["ans"] = [float(Fraction_eval(v["_expr"])) if isinstance(v, dict) else v for v in sub["ans"]]
    assert len(dgv) == 15, len(dgv)
    json.dump({"source": "blas/testblas/dgemv.go:34-629 (DgemvCases; harness :631-683, dSliceTolEqual 1e-14 normalised)", "cases": dgv},
              open(os.path.join(OUT, "dgemv_cases.json"), "w"), indent=1)
    red = extract_reduce_cases(os.path.join(REF, "reduce_test.go"))
    assert [len(red[k]) for k in ("testSum", "testMax", "testMin", "testArgMax", "testArgMin", "testSumSquares")] == [5, 5, 5, 6, 6, 5], red
    json.dump({"source": "reduce_test.go:28-240 (tables of testSum/testMax/testMin/testArgMax/testArgMin/testSumSquares; floatEquals tol 1e-5)",
               "cases": red}, open(os.path.join(OUT, "reduce_cases.json"), "w"), indent=1)
    print("dgemm cases:", len(dg), "one-vector:", len(keep1), "two-vector:", len(keep2), "gelu points:", len(pts))


if __name__ == "__main__":
    sys.exit(main())
