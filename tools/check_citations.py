"""Checks every `<reference file>:<lines>` citation in the repo against /root/reference (this container only; the GPU box
has no reference tree).  A cited range must exist in the file, lie past the 19-line licence header and contain at least
one line of code (not only scaladoc / blank lines) -- the mistakes a reviewer following a citation would land on.

    python tools/check_citations.py            # prints the offenders, exit status 1 if any
"""
import os
import re
import sys

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ALIASES = {
    "S/": "flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/",
    "T/": "flink-jpmml-scala/src/test/scala/io/radicalbit/flink/pmml/scala/",
    "E/": "flink-jpmml-examples/src/main/scala/io/radicalbit/examples/",
    "A/": "flink-jpmml-assets/src/main/resources/",
}
SCAN = ["include", "flink-jpmml_b200", "integration", "oracle", "tests", "tools", "bench.py", "__graft_entry__.py", "DESIGN.md", "INTEGRATION.md", "BASELINE.md", "README.md"]
SKIP_DIRS = {"build", "__pycache__", "golden", "_ref"}
CITE = re.compile(r"((?:[STEA]/)?[A-Za-z0-9_./-]*[A-Za-z0-9_]+\.(?:scala|xml|sbt|md)):(\d+(?:-\d+)?(?:, ?\d+(?:-\d+)?)*)")


def all_reference_files():
    out = []
    for d, _, fs in os.walk(REF):
        for f in fs:
            out.append(os.path.relpath(os.path.join(d, f), REF))
    return out


def resolve(path, files):
    for a, full in ALIASES.items():
        if path.startswith(a):
            cand = full + path[len(a):]
            return cand if cand in files else None
    hits = [f for f in files if f == path or f.endswith("/" + path)]
    return hits[0] if len(hits) == 1 else None     # bare or ambiguous names (package.scala) cannot be checked


def is_code(line):
    t = line.strip()
    return bool(t) and not t.startswith(("*", "/**", "*/", "//", "/*"))


def main():
    if not os.path.isdir(REF):
        print("no reference tree here: nothing to check"); return 0
    files = set(all_reference_files()); bad = []; checked = 0
    for entry in SCAN:
        p = os.path.join(ROOT, entry)
        walk = [(os.path.dirname(p), [], [os.path.basename(p)])] if os.path.isfile(p) else os.walk(p)
        for d, dirs, fs in walk:
            dirs[:] = [x for x in dirs if x not in SKIP_DIRS]
            for f in fs:
                if not f.endswith((".h", ".hpp", ".cc", ".cu", ".cuh", ".py", ".md", ".scala", ".c")) or f == "check_citations.py":
                    continue
                src = os.path.join(d, f)
                for ln, line in enumerate(open(src, errors="replace"), 1):
                    for m in CITE.finditer(line):
                        target = resolve(m.group(1), files)
                        if target is None:
                            continue
                        text = open(os.path.join(REF, target), errors="replace").read().split("\n")
                        header = 19 if target.endswith((".scala", ".sbt")) else 0
                        for rng in re.split(r", ?", m.group(2)):
                            a, _, b = rng.partition("-"); a = int(a); b = int(b or a); checked += 1
                            why = None
                            if b < a or b > len(text): why = f"outside the file ({len(text)} lines)"
                            elif a <= header: why = "inside the licence header"
                            elif target.endswith(".scala") and not any(is_code(x) for x in text[a - 1:b]): why = "only comments / blank lines"
                            if why: bad.append(f"{os.path.relpath(src, ROOT)}:{ln}: {m.group(1)}:{rng} -> {why}")
    for b in bad: print(b)
    print(f"{checked} cited ranges checked, {len(bad)} bad")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
