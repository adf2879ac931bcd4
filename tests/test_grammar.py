"""GBNF grammar matcher (csrc/grammar.cc) through the C ABI, on the host -- no device needed.

Anchors: the grammar texts the reference's own tests expect its JSON-schema converter to emit
(pkg/functions/grammars/json_schema_test.go, llama31_schema_test.go; collected by tools/make_grammar_goldens.py into
tests/golden/grammars.json), known-answer cases, and the Python restatement in oracle/grammar_ref.py."""
import json
import os

import numpy as np
import pytest

from localai_b200 import capi
from oracle import grammar_ref as G

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "grammars.json")))

CALL_EVENT = '{"arguments": {"date": "2024-01-01", "time": "10:00", "title": "café \\"x\\" \\u00e9 日本"}, "function": "create_event"}'
CALL_SEARCH = '{"arguments":{"query":"b200"},"function":"search"}'


def both(gbnf, text):
    """(bytes accepted, may stop) from the C++ matcher, checked against the Python restatement"""
    nb, stop = capi.grammar_match(gbnf, text)
    n_cp, stop_ref = G.match(G.parse(gbnf), text)
    assert nb == len(text[:n_cp].encode("utf-8")) and stop == stop_ref, (text, nb, n_cp, stop, stop_ref)
    return nb, stop


@pytest.mark.parametrize("name", sorted(GOLDEN))
def test_reference_grammars_parse_and_accept_their_calls(name):
    g = GOLDEN[name]
    key = "name" if '\\"name\\"' in g else "function"
    calls = [c.replace("function", key) for c in (CALL_EVENT, CALL_SEARCH)]
    if name.startswith("testllama31"):
        calls = ['<function=create_event>{{"date":"d","time":"t","title":"x"}}</function>', '<function=search>{{"query": "q" }}</function>']
    for c in calls:
        # the root rules differ in how a call is wrapped (bare, behind "suffix", inside an array): one form must be a sentence
        forms = [c, "suffix" + c, "suffix[\n" + c + ",\n" + c + "]", "[\n" + c + "]"]
        done = [f for f in forms if both(g, f) == (len(f.encode()), True)]
        assert done, (name, [both(g, f) for f in forms])
        c = done[0]
        # every proper prefix is viable; a mutation dies where it was made
        for cut in (1, len(c) // 2, len(c) - 1):
            assert both(g, c[:cut])[0] == len(c[:cut].encode())
        bad = c[:1] + "\x01" + c[1:]
        nb, _ = both(g, bad)
        assert nb <= 1 or name.startswith("rootResult")  # those roots also take free text up to the first quote
    both(g, "plain text, not a call")
    both(g, "")


def test_reference_grammars_random_mutations_agree_with_restatement():
    rng = np.random.default_rng(5)
    alphabet = list('{}[]":, \\abcdefu0123456789é\n')
    for name in ("inputResult1", "inputResult4", "rootResult3", "rootResult5"):
        g = GOLDEN[name]
        key = "name" if '\\"name\\"' in g else "function"
        base = CALL_EVENT.replace("function", key)
        for _ in range(60):
            s = list(base)
            for _ in range(int(rng.integers(1, 4))):
                i = int(rng.integers(0, len(s)))
                op = int(rng.integers(0, 3))
                if op == 0:
                    s[i] = alphabet[int(rng.integers(0, len(alphabet)))]
                elif op == 1:
                    del s[i]
                else:
                    s.insert(i, alphabet[int(rng.integers(0, len(alphabet)))])
            both(g, "".join(s))


KNOWN = [
    # grammar, text, bytes accepted, may stop
    ('root ::= "a"{2,4}\n', "a", 1, False),
    ('root ::= "a"{2,4}\n', "aaa", 3, True),
    ('root ::= "a"{2,4}\n', "aaaaa", 4, True),
    ('root ::= "a"{3}\n', "aaaa", 3, True),
    ('root ::= "a"{2,} "b"\n', "aaaaaab", 7, True),
    ('root ::= "a"{0,2} "b"\n', "b", 1, True),
    ('root ::= ("ab" | "cd")* "e"?\n', "abcdab", 6, True),
    ('root ::= ("ab" | "cd")* "e"?\n', "abce", 3, False),  # "c" opens "cd"; the count is of viable bytes
    ('root ::= [^\\n]+ "\\n"\n', "héllo\n", 7, True),
    ('root ::= . . .\n', "日本語!", 9, True),
    ('root ::= [a-cx-z\\]]+\n', "abz]d", 4, True),
    ('root ::= "\\x41\\u00e9\\U0001F600"\n', "Aé😀", 7, True),
    ('# comment\nroot ::= item+ # trailing\nitem ::= "-" [0-9]+ "\\n"\n', "-1\n-22\n", 7, True),
    ('root ::= a | b\na ::= "x" a | "x"\nb ::= "y"\n', "xxx", 3, True),
    ('root\t::=\n  "a" |\n  "b"\n', "b", 1, True),  # a line may end in "|", the next one continues the rule
]


@pytest.mark.parametrize("case", range(len(KNOWN)))
def test_known_answers(case):
    g, text, nb, stop = KNOWN[case]
    assert both(g, text) == (nb, stop)


@pytest.mark.parametrize("bad", [
    'root ::= foo\n',                      # undefined rule
    'foo ::= "a"\n',                       # no root
    'root ::= root "a" | "b"\n',           # direct left recursion
    'root ::= a\na ::= b "x" | "y"\nb ::= a?\n',  # left recursion through a nullable rule
    'root ::= "a\n',                       # literal not closed
    'root ::= ("a" | "b"\n',               # group not closed
    'root ::= [a-\n',                      # class not closed
    'root ::= "a"{2,1}\n',                 # bounds
    'root ::= "a"{x}\n',
    'root ::= "\\q"\n',                    # unknown escape
    'root ::= "\\x4"\n',                   # short hex escape
    'root = "a"\n',                        # no ::=
    'root ::= "a" "b"\n)',                 # junk after the last rule
    'root ::=\n  "a"\n  | "b"\n',          # outside a group a rule ends at the newline: "|" cannot open a line
    'root ::= *\n',                        # nothing to repeat
    'root ::= "a"**\n',                    # a second operator has no item of its own
])
def test_parse_errors(bad):
    with pytest.raises(RuntimeError, match="grammar"):
        capi.grammar_match(bad, "a")
    with pytest.raises(G.GrammarError):
        G.match(G.parse(bad), "a")


def test_token_mask_known_answers_incl_partial_utf8():
    g = 'root ::= "{" space "\\"a\\"" space ":" space num "}" space\nnum ::= [0-9]+ ("." [0-9]{1,3})?\nspace ::= " "?\n'
    pieces = [b"2", b".5", b"}", b"} ", b"}}", b"x", b"", b"<eos>", b"\xc3"]
    m = capi.grammar_token_mask(g, '{"a":1', pieces, [0, 0, 0, 0, 0, 0, 0, 1, 0])
    assert m.tolist() == [True, True, True, True, False, False, False, False, False]
    m = capi.grammar_token_mask(g, '{"a":1}', [b"2", b" ", b"<eos>", b"  "], [0, 0, 1, 0])
    assert m.tolist() == [False, True, True, False]  # complete: only the optional space or the end
    # pieces that end, or begin, inside a multi-byte sequence (byte-fallback tokens)
    g2 = 'root ::= [^a-z]* "é" [é-ü]\n'
    m = capi.grammar_token_mask(g2, "", [b"\xc3", b"\xc3\xa9", b"A\xc3", b"a", b"\xe2\x82", b"\xa9", b"\xf0\x9f"])
    assert m.tolist() == [True, True, True, False, True, False, True]
    m = capi.grammar_token_mask(g2, b"\xc3", [b"\xa9", b"\xa9\xc3", b"a", b"\xc3\xa9", b"\xa9\xc3\xbc", b"\xa9\xc3\xbf"])
    assert m.tolist() == [True, True, False, False, True, True]  # last: é then ÿ, which [^a-z]* still takes
    g3 = 'root ::= "é" [é-ü]\n'
    m = capi.grammar_token_mask(g3, b"\xc3", [b"\xa9\xc3\xbc", b"\xa9\xc3\xbf", b"\xa9\xc3", b"\xa9\xc2", b"\xa9\xe2"])
    assert m.tolist() == [True, False, True, False, False]  # a tail stands for the range of code points it can become


def test_token_mask_equals_viable_prefix_definition():
    """bit t  <=>  accepted + piece[t] is a viable prefix -- over a vocabulary of real substrings and junk"""
    rng = np.random.default_rng(11)
    g = GOLDEN["inputResult1"]
    rules = G.parse(g)
    pieces = set()
    for s in (CALL_EVENT, CALL_SEARCH, 'xyz "\\ {}[]:, 0123 \\u00e9 \\n'):
        for _ in range(400):
            i = int(rng.integers(0, len(s)))
            pieces.add(s[i:i + int(rng.integers(1, 7))])
    pieces = sorted(pieces)
    for cut in (0, 1, 15, 30, 47, 80, len(CALL_EVENT) - 1, len(CALL_EVENT)):
        acc = CALL_EVENT[:cut]
        eog = [0] * len(pieces) + [1]
        m = capi.grammar_token_mask(g, acc, [p.encode() for p in pieces] + [b"</s>"], eog)
        want = [G.viable(rules, acc + p) for p in pieces] + [G.match(rules, acc)[1]]
        assert m.tolist() == want, cut


def test_sentences_agree_with_xgrammar():
    """An independent GBNF implementation (xgrammar, the grammar engine vLLM ships; it reads llama.cpp's grammar format)
    decides "is this text a complete sentence of the grammar" like this engine's matcher: on the reference's own grammar
    texts with calls, wrapped calls, prefixes and random mutations of them, and on the known-answer grammars above."""
    xg = pytest.importorskip("xgrammar")
    from xgrammar.testing import _is_grammar_accept_string as accepts

    rng = np.random.default_rng(9)
    alphabet = list('{}[]":, \\abcdefu0123456789é\n')
    n_checked = n_sentences = 0
    for name in sorted(GOLDEN):
        g = GOLDEN[name]
        theirs = xg.Grammar.from_ebnf(g, root_rule_name="root")
        key = "name" if '\\"name\\"' in g else "function"
        calls = [c.replace("function", key) for c in (CALL_EVENT, CALL_SEARCH)]
        if name.startswith("testllama31"):
            calls = ['<function=create_event>{{"date":"d","time":"t","title":"x"}}</function>', '<function=search>{{"query": "q" }}</function>']
        texts = []
        for c in calls:
            texts += [c, "suffix" + c, "suffix[\n" + c + ",\n" + c + "]", "[\n" + c + "]", c[:-1], c[:len(c) // 2], "plain text", ""]
            for _ in range(25):
                s = list(c)
                for _ in range(int(rng.integers(1, 3))):
                    i = int(rng.integers(0, len(s)))
                    op = int(rng.integers(0, 3))
                    if op == 0:
                        s[i] = alphabet[int(rng.integers(0, len(alphabet)))]
                    elif op == 1:
                        del s[i]
                    else:
                        s.insert(i, alphabet[int(rng.integers(0, len(alphabet)))])
                texts.append("".join(s))
        for t in texts:
            nb, stop = capi.grammar_match(g, t)
            mine = bool(stop) and nb == len(t.encode("utf-8"))
            assert mine == bool(accepts(theirs, t)), (name, t, nb, stop)
            n_checked += 1
            n_sentences += mine
    for gbnf, text, nb_want, stop_want in KNOWN:
        try:
            theirs = xg.Grammar.from_ebnf(gbnf, root_rule_name="root")
        except Exception:  # a construct xgrammar does not read
            continue
        for t in (text, text[:-1], text + "a", text[:nb_want]):
            nb, stop = capi.grammar_match(gbnf, t)
            mine = bool(stop) and nb == len(t.encode("utf-8"))
            assert mine == bool(accepts(theirs, t)), (gbnf, t, nb, stop)
            n_checked += 1
            n_sentences += mine
    assert n_checked > 600 and 40 < n_sentences < n_checked - 200  # both answers occur often enough to mean something


def test_token_masks_agree_with_xgrammar():
    """... and the vocabulary mask after an accepted prefix (what constrains the sampler) is xgrammar's next-token bitmask,
    over a vocabulary of real substrings, junk and an end-of-generation token."""
    xg = pytest.importorskip("xgrammar")
    rng = np.random.default_rng(11)
    pieces = set()
    for s in (CALL_EVENT, CALL_SEARCH, 'xyz "\\ {}[]:, 0123 \\u00e9 \\n', "suffix[\n"):
        for _ in range(400):
            i = int(rng.integers(0, len(s)))
            pieces.add(s[i:i + int(rng.integers(1, 7))])
    pieces = sorted(p for p in pieces if p)
    vocab = [p.encode() for p in pieces] + [b"</s>"]
    ti = xg.TokenizerInfo(vocab, vocab_type=xg.VocabType.RAW, stop_token_ids=[len(vocab) - 1])
    compiler = xg.GrammarCompiler(ti)
    idx = np.arange(len(vocab))
    n_masks = 0
    for name in ("inputResult1", "inputResult2", "inputResult4", "rootResult0", "rootResult3", "rootResult5"):
        g = GOLDEN[name]
        compiled = compiler.compile_grammar(xg.Grammar.from_ebnf(g, root_rule_name="root"))
        key = "name" if '\\"name\\"' in g else "function"
        call = CALL_EVENT.replace("function", key)
        forms = [f for f in (call, "suffix" + call, "[\n" + call + "]") if capi.grammar_match(g, f) == (len(f.encode()), True)]
        assert forms, name
        text = forms[0]
        for cut in sorted({0, 1, 2, 15, 30, 47, 80, len(text) // 2, len(text) - 1, len(text)}):
            acc = text[:cut]
            m = xg.GrammarMatcher(compiled)
            assert not acc or m.accept_string(acc)
            bm = xg.allocate_token_bitmask(1, ti.vocab_size)
            m.fill_next_token_bitmask(bm)
            bits = bm.numpy()[0].view(np.uint32)
            theirs = ((bits[idx >> 5] >> (idx & 31).astype(np.uint32)) & 1).astype(bool)
            mine = capi.grammar_token_mask(g, acc, vocab, [0] * len(pieces) + [1])
            assert mine.tolist() == theirs.tolist(), (name, cut, [vocab[i] for i in np.nonzero(mine != theirs)[0][:5]])
            n_masks += 1
    assert n_masks >= 50
