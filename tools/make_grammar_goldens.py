#!/usr/bin/env python3
"""Collects the GBNF texts the reference's own tests expect its JSON-schema converter to emit
(pkg/functions/grammars/json_schema_test.go, llama31_schema_test.go) into tests/golden/grammars.json.
They are the grammar strings LocalAI sends in PredictOptions.Grammar for tools / function calls, so the matcher must
parse every one of them.  Run in the container that has /root/reference; the JSON travels, the Go files do not.
"""
import json
import os
import re

REF = "/root/reference/pkg/functions/grammars"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "grammars.json")


def main():
    out = {}
    src = open(os.path.join(REF, "json_schema_test.go")).read()
    for m in re.finditer(r"^\t(inputResult\d+) = `(.*?)`", src, re.S | re.M):
        out[m.group(1)] = m.group(2)
    # rootResult(s) is a template: everything between "return `" and the closing "`", with the root rule spliced in
    t = re.search(r"func rootResult\(s string\) string \{\n\treturn `(.*?)` \+ s \+ `(.*?)`\n\}", src, re.S)
    for i, root in enumerate(re.findall(r"rootResult\(`(.*?)`\)", src)):
        # every test that uses the template joins this line to it (json_schema_test.go:284-399)
        out[f"rootResult{i}"] = t.group(1) + root + t.group(2) + "\nmixedstring ::= freestring | freestring arr | freestring realvalue"
    src = open(os.path.join(REF, "llama31_schema_test.go")).read()
    for m in re.finditer(r"^\t(testllama31inputResult\d+) = `(.*?)`", src, re.S | re.M):
        out[m.group(1)] = m.group(2)
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print(f"{len(out)} grammars -> {OUT}")


if __name__ == "__main__":
    main()
