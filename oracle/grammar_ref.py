"""TEST INFRASTRUCTURE ONLY -- a small Python restatement of GBNF matching (llama.cpp's llama-grammar.cpp at the pin
SURVEY.md 8c names, d5cb868; not vendored in the reference tree, reached through gpt_sampler_init / _sample / _accept at
backend/cpp/llama/grpc-server.cpp:890,1946,1948).  The product's matcher is localai_b200/csrc/grammar.cc; this file
checks it and is never imported by the product.

Parity: UNPINNED against upstream binaries (llama.cpp is absent here).  It is anchored on the grammar texts the
reference's own tests expect its JSON-schema converter to emit (pkg/functions/grammars/json_schema_test.go:60-84,
llama31_schema_test.go) -- they must parse, accept the function-call JSON they were written for and reject its
mutations -- and on known-answer cases in tests/test_grammar.py.

Representation differs from the C++ on purpose: a state is a set of continuations, a continuation is a tuple of
"rest of sequence" tuples (innermost last); items are ("c", negated, ranges) | ("." ,) | ("r", rule name).
"""
from __future__ import annotations


class GrammarError(ValueError):
    pass


class _P:
    def __init__(self, text: str):
        self.s, self.i, self.rules, self.n_sym, self.order = text, 0, {}, 0, []

    def peek(self, k=0):
        return self.s[self.i + k] if self.i + k < len(self.s) else ""

    def space(self, nl):
        while self.peek():
            c = self.peek()
            if c == "#":
                while self.peek() and self.peek() not in "\r\n":
                    self.i += 1
            elif c in " \t" or (nl and c in "\r\n"):
                self.i += 1
            else:
                break

    @staticmethod
    def word(c):
        return c != "" and (c.isascii() and (c.isalnum() or c == "-"))

    def name(self):
        j = self.i
        while self.word(self.peek()):
            self.i += 1
        if j == self.i:
            raise GrammarError(f"expecting a name at {self.i}")
        return self.s[j:self.i]

    def sym(self, name):
        if name not in self.order:
            self.order.append(name)
        return name

    def fresh(self, base):
        return self.sym(f"{base}_{len(self.order)}")

    def char(self):
        c = self.peek()
        if c == "":
            raise GrammarError("unexpected end")
        if c == "\\":
            e = self.peek(1)
            self.i += 2
            n = {"x": 2, "u": 4, "U": 8}.get(e)
            if n:
                h = self.s[self.i:self.i + n]
                if len(h) != n or any(ch not in "0123456789abcdefABCDEF" for ch in h):
                    raise GrammarError("bad hex escape")
                self.i += n
                return int(h, 16)
            if e in "\\\"[]":
                return ord(e)
            if e in "trn":
                return ord({"t": "\t", "r": "\r", "n": "\n"}[e])
            raise GrammarError("unknown escape")
        self.i += 1
        return ord(c)

    def integer(self):
        j = self.i
        while self.peek().isdigit() and self.peek().isascii():
            self.i += 1
        if j == self.i:
            raise GrammarError("expecting an integer")
        return int(self.s[j:self.i])

    def repeat(self, seq, start, rule, lo, hi):
        if start >= len(seq):
            raise GrammarError("nothing to repeat")
        symb = seq[start:]
        if lo == 0:
            del seq[start:]
        for _ in range(1, lo):
            seq.extend(symb)
        n_opt = 1 if hi < 0 else hi - lo
        last = None
        for i in range(n_opt):
            nm = self.fresh(rule)
            body = list(symb)
            if hi < 0:
                body.append(("r", nm))
            elif i > 0:
                body.append(("r", last))
            self.rules[nm] = [tuple(body), ()]
            last = nm
        if n_opt > 0:
            seq.append(("r", last))

    def sequence(self, rule, nested):
        seq, start = [], 0
        while self.peek():
            c = self.peek()
            if c == '"':
                self.i += 1
                start = len(seq)
                while self.peek() != '"':
                    cp = self.char()
                    seq.append(("c", False, ((cp, cp),)))
                self.i += 1
                self.space(nested)
            elif c == "[":
                self.i += 1
                neg = self.peek() == "^"
                self.i += neg
                rs = []
                while self.peek() != "]":
                    lo = self.char()
                    hi = lo
                    if self.peek() == "-" and self.peek(1) != "]":
                        self.i += 1
                        hi = self.char()
                    rs.append((lo, hi))
                self.i += 1
                start = len(seq)
                seq.append(("c", neg, tuple(rs)))
                self.space(nested)
            elif self.word(c):
                nm = self.sym(self.name())
                self.space(nested)
                start = len(seq)
                seq.append(("r", nm))
            elif c == "(":
                self.i += 1
                self.space(True)
                nm = self.fresh(rule)
                self.alternates(rule, nm, True)
                start = len(seq)
                seq.append(("r", nm))
                if self.peek() != ")":
                    raise GrammarError("expecting )")
                self.i += 1
                self.space(nested)
            elif c == ".":
                start = len(seq)
                seq.append((".",))
                self.i += 1
                self.space(nested)
            elif c in "*+?":
                self.i += 1
                self.space(nested)
                self.repeat(seq, start, rule, 1 if c == "+" else 0, 1 if c == "?" else -1)
                start = len(seq)
            elif c == "{":
                self.i += 1
                self.space(nested)
                lo = hi = self.integer()
                self.space(nested)
                if self.peek() == ",":
                    self.i += 1
                    self.space(nested)
                    if self.peek().isdigit():
                        hi = self.integer()
                        self.space(nested)
                    else:
                        hi = -1
                if self.peek() != "}":
                    raise GrammarError("expecting }")
                self.i += 1
                if 0 <= hi < lo:
                    raise GrammarError("bad repetition bounds")
                self.space(nested)
                self.repeat(seq, start, rule, lo, hi)
                start = len(seq)
            else:
                break
        return tuple(seq)

    def alternates(self, rule, target, nested):
        alts = [self.sequence(rule, nested)]
        while self.peek() == "|":
            self.i += 1
            self.space(True)
            alts.append(self.sequence(rule, nested))
        self.rules[target] = alts

    def parse(self):
        self.space(True)
        while self.peek():
            nm = self.name()
            self.space(False)
            self.sym(nm)
            if self.s[self.i:self.i + 3] != "::=":
                raise GrammarError("expecting ::=")
            self.i += 3
            self.space(True)
            self.alternates(nm, nm, False)
            if self.peek() == "\r":
                self.i += 2 if self.peek(1) == "\n" else 1
            elif self.peek() == "\n":
                self.i += 1
            elif self.peek():
                raise GrammarError(f"expecting newline at {self.i}: {self.s[self.i:self.i + 20]!r}")
            self.space(True)
        if "root" not in self.rules:
            raise GrammarError("no root rule")
        for nm in self.order:
            if nm not in self.rules:
                raise GrammarError(f"undefined rule {nm}")
        return self.rules


def parse(text: str) -> dict:
    return _P(text).parse()


def _expand(rules, cont, out, depth=0):
    """cont: tuple of sequences (innermost last).  Collect continuations whose next item is a terminal (or () = done)."""
    if depth > 400:
        raise GrammarError("left recursion")
    while cont and not cont[-1]:
        cont = cont[:-1]
    if not cont:
        out.add(())
        return
    item = cont[-1][0]
    if item[0] != "r":
        out.add(cont)
        return
    rest = cont[:-1] + (cont[-1][1:],)
    for alt in rules[item[1]]:
        _expand(rules, rest + (alt,), out, depth + 1)


def _matches(item, cp):
    if item[0] == ".":
        return True
    _, neg, ranges = item
    return any(lo <= cp <= hi for lo, hi in ranges) != neg


def start(rules):
    out = set()
    for alt in rules["root"]:
        _expand(rules, (alt,), out)
    return out


def advance(rules, state, cp):
    nxt = set()
    for cont in state:
        if cont and _matches(cont[-1][0], cp):
            _expand(rules, cont[:-1] + (cont[-1][1:],), nxt)
    return nxt


def match(rules, text: str):
    """-> (code points of `text` accepted before the first that does not fit, may the grammar stop there)."""
    state = start(rules)
    n = 0
    for ch in text:
        nxt = advance(rules, state, ord(ch))
        if not nxt:
            break
        state, n = nxt, n + 1
    return n, () in state


def viable(rules, text: str) -> bool:
    return match(rules, text)[0] == len(text)
