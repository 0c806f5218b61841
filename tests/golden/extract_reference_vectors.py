#!/usr/bin/env python
"""Extracts literal golden vectors from the reference's own Go tests into JSON fixtures.

Run in the build container (needs /root/reference, which does not exist on the GPU box):
    python tests/golden/extract_reference_vectors.py
Only test *data* (table rows of names / integers) is extracted, never code: each Go
``tests := []struct{...}{ {...}, ... }`` literal is parsed with a small composite-literal
parser and written as JSON; the checks themselves are restated in tests/test_oracle_*.py.

Outputs (tests/golden/)
  container_combinations.json  roaring/roaring_internal_test.go TestContainerCombinations (:2713-3345)
  roaring_tables.json          every table-driven test listed in TABLES below, keyed by Go test name
"""
import json
import os
import re

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))

CONSTS = {"maxContainerVal": 0xFFFF, "bitmapN": 1024, "true": True, "false": False, "nil": None,
          "ArrayMaxSize": 4096, "runMaxSize": 2048, "containerArray": 1, "containerBitmap": 2, "containerRun": 3,
          "containerNil": 0, "maxBitmap": 0xFFFFFFFFFFFFFFFF}


class P:
    """Recursive-descent parser for a Go composite literal (data only)."""

    def __init__(self, s, i=0):
        self.s, self.i = s, i

    def ws(self):
        while True:
            while self.i < len(self.s) and self.s[self.i] in " \t\r\n":
                self.i += 1
            if self.s.startswith("//", self.i):
                self.i = self.s.index("\n", self.i)
                continue
            if self.s.startswith("/*", self.i):
                self.i = self.s.index("*/", self.i) + 2
                continue
            return

    def peek(self):
        self.ws()
        return self.s[self.i] if self.i < len(self.s) else ""

    def composite(self):
        """'{' elem, elem ... '}' -> dict (keyed) or list."""
        assert self.peek() == "{", self.s[self.i:self.i + 40]
        self.i += 1
        items, keyed = [], {}
        while True:
            if self.peek() == "}":
                self.i += 1
                break
            save = self.i
            m = re.match(r"([A-Za-z_]\w*)\s*:", self.s[self.i:])
            if m:
                self.i += m.end()
                keyed[m.group(1)] = self.value()
            else:
                self.i = save
                items.append(self.value())
            if self.peek() == ",":
                self.i += 1
        return keyed if keyed else items

    def value(self):
        c = self.peek()
        if c == "{":
            return self.composite()
        if c == '"':
            j = self.s.index('"', self.i + 1)
            v = self.s[self.i + 1:j]
            self.i = j + 1
            return v
        # type-prefixed literal / conversion / call
        m = re.match(r"((?:\[\])*[A-Za-z_][\w.]*)\s*([{(])", self.s[self.i:])
        if m and not re.match(r"(uint64|uint16|int32|int64|uint32|int|uint)\(", self.s[self.i:]):
            name, br = m.group(1), m.group(2)
            if br == "{":
                self.i += m.end() - 1
                return self.composite()
            # call or conversion: NAME(args)
            self.i += m.end()
            args = []
            while self.peek() != ")":
                args.append(self.value())
                if self.peek() == ",":
                    self.i += 1
            self.i += 1
            if name.startswith("[]"):
                return args[0] if args else None  # []uint16(nil)
            return {"call": name, "args": args}
        # scalar expression up to , } )
        j, depth = self.i, 0
        while j < len(self.s):
            ch = self.s[j]
            if ch in "([":
                depth += 1
            elif ch in ")]":
                if depth == 0:
                    break
                depth -= 1
            elif ch in ",}" and depth == 0:
                break
            j += 1
        expr = self.s[self.i:j].strip()
        self.i = j
        expr_py = re.sub(r"\b(uint64|uint16|int32|int64|uint32|int|uint)\(", "(", expr)
        try:
            return eval(expr_py, {"__builtins__": {}}, dict(CONSTS))
        except Exception:
            return {"expr": expr}


def table(src, func, marker=r":= \[\]struct\s*\{"):
    start = src.index("func " + func + "(")
    m = re.compile(marker).search(src, start)
    # skip the struct type body, then parse the literal list that follows
    p = P(src, m.end() - 1)
    depth, i = 0, p.i
    while True:  # match the struct type braces
        if src[i] == "{":
            depth += 1
        elif src[i] == "}":
            depth -= 1
            if depth == 0:
                break
        i += 1
    p.i = i + 1
    return p.composite()


INTERNAL_TABLES = [
    "TestRunAppendInterval", "TestContainerRunAdd", "TestBitmapCountRange", "TestIntersectionCountArrayBitmap2",
    "TestRunRemove", "TestIntersectionCountRunRun", "TestIntersectArrayRun", "TestIntersectRunRun",
    "TestIntersectBitmapRunBitmap", "TestIntersectBitmapRunArray", "TestUnionRunRun", "TestUnionArrayRun",
    "TestBitmapSetRange", "TestArrayToBitmap", "TestBitmapToArray", "TestRunToBitmap", "TestBitmapToRun",
    "TestArrayToRun", "TestRunToArray", "TestBitmapZeroRange", "TestUnionBitmapRun", "TestBitmapCountRuns",
    "TestArrayCountRuns", "TestDifferenceArrayRun", "TestDifferenceRunArray", "TestDifferenceRunBitmap",
    "TestDifferenceBitmapRun", "TestDifferenceBitmapArray", "TestDifferenceBitmapBitmap", "TestDifferenceRunRun",
    "TestXorArrayRun", "TestXorRunRun", "TestBitmapXorRange", "TestXorBitmapRun", "TestIntersectArrayBitmap",
    "TestShiftArray", "TestShiftBitmap", "TestShiftRun", "TestRunBinSearchContains",
]


def container_combinations():
    src = open(os.path.join(REF, "roaring/roaring_internal_test.go")).read()
    start = src.index("func TestContainerCombinations")
    body = src[start: src.index("for _, testOp := range testOps", start)]
    rows = [list(m.groups()) for m in
            re.finditer(r'^\s*\{(intersect|union|difference|xor|flip), "(\w*)", "(\w*)", "(\w*)"\},', body, re.M)]
    assert len(rows) > 300, len(rows)
    with open(os.path.join(OUT, "container_combinations.json"), "w") as f:
        f.write('{"source": "roaring/roaring_internal_test.go:2713-3345 TestContainerCombinations",\n"rows": [\n')
        f.write(",\n".join(json.dumps(r) for r in rows))
        f.write("\n]}\n")
    print("container_combinations:", len(rows), "rows")


def roaring_tables():
    src = open(os.path.join(REF, "roaring/roaring_internal_test.go")).read()
    out = {}
    for name in INTERNAL_TABLES:
        try:
            out[name] = table(src, name)
        except Exception as e:  # noqa
            print("  !! could not extract", name, repr(e)[:80])
    with open(os.path.join(OUT, "roaring_tables.json"), "w") as f:
        f.write('{"source": "roaring/roaring_internal_test.go (table-driven tests, literal rows only)"')
        for k, v in out.items():
            f.write(",\n" + json.dumps(k) + ": [\n" + ",\n".join(json.dumps(r) for r in v) + "\n]")
        f.write("\n}\n")
    print("roaring_tables:", {k: len(v) for k, v in out.items()})


def official_roaring():
    """TestUnmarshalRoaringWithNoErrors / WithErrors (roaring/roaring_internal_test.go:3359-3435): the hex
    literals, the expected counts / error texts, and the bytes of the external fixture the test reads."""
    src = open(os.path.join(REF, "roaring/roaring_internal_test.go")).read()
    body = src[src.index("func TestUnmarshalRoaringWithNoErrors"):src.index("func BenchmarkUnionBitmapBitmapInPlace")]
    ok = []
    for m in re.finditer(r'roaringData:\s*"([0-9A-Fa-f]+)",\s*count:\s*(\d+),\s*expectedBits:\s*"\[([^\]]*)\]"', body):
        ok.append({"hex": m.group(1), "count": int(m.group(2)), "bits": [int(x) for x in m.group(3).split()]})
    m = re.search(r'roaringFileName:\s*"([^"]+)",\s*count:\s*(\d+)', body)
    blob = open(os.path.join(REF, "roaring", m.group(1)), "rb").read()
    ok.append({"hex": blob.hex(), "count": int(m.group(2)), "bits": None, "file": m.group(1)})
    bad = []
    for m in re.finditer(r'hexString:\s*"([0-9A-Fa-f]+)",(?:\s*expectedError:\s*"([^"]*)",)?', body):
        bad.append({"hex": m.group(1), "error": m.group(2)})
    with open(os.path.join(OUT, "official_roaring.json"), "w") as f:
        json.dump({"source": "roaring/roaring_internal_test.go:3359-3435", "ok": ok, "errors": bad}, f, indent=0)
    print("official_roaring.json:", len(ok), "valid images,", len(bad), "header cases")


def time_views():
    """TestViewsByTimeRange (time_internal_test.go:87-160): (start, end, quantum) -> view names."""
    src = open(os.path.join(REF, "time_internal_test.go")).read()
    body = src[src.index("func TestViewsByTimeRange"):]
    body = body[:body.index("\nfunc ", 10)]
    cases = []
    pat = re.compile(r'viewsByTimeRange\("F", mustParseTime\("([^"]+)"\), mustParseTime\("([^"]+)"\), mustParseTimeQuantum\("([A-Z]+)"\)\)\s*'
                     r'if !reflect.DeepEqual\(a, \[\]string\{([^}]*)\}\)', re.S)
    for m in pat.finditer(body):
        cases.append({"start": m.group(1), "end": m.group(2), "quantum": m.group(3), "views": re.findall(r'"([^"]+)"', m.group(4))})
    with open(os.path.join(OUT, "time_views.json"), "w") as f:
        json.dump({"source": "time_internal_test.go:87-160 TestViewsByTimeRange", "cases": cases}, f, indent=0)
    print("time_views.json:", len(cases), "cases")


if __name__ == "__main__":
    time_views()
    official_roaring()
    container_combinations()
    roaring_tables()
