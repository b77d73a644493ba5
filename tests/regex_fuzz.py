"""Grammar-based generator of regexes inside the SPEC.md section 3 subset, and of lines that
exercise them (TEST INFRASTRUCTURE: used by the CPU three-way differential test and by
tests/soak_parity.py on the GPU).

A regex is drawn as a small tree (literal runs, classes, escapes, groups, alternations, flags,
quantifiers, anchors at the ends); every tree can also *sample* a string it matches, so that the
lines of a test are not only random noise: for every regex there are exact matches, matches
embedded in other text, and one-character mutations of matches (the near misses a translator
bug would get wrong).
"""
from __future__ import annotations

import random
from typing import List, Tuple

LETTERS = "abcxyzABZ"
WORDS = ["error", "OOM", "exit code", "timeout", "refused", "killed", "at ", "panic:", "E12", "db-7", "a.b", "x+y", "[warn]", "été", "€5"]
PUNCT = " -_.:/[]()+*?$^|{}\\\"'#@=,;<>!~&%"
NOISE = "abcxyzABZ019 _-.:/[]()\t€é+*$^|\\\"'#"


def _esc_lit(ch: str, rnd: random.Random) -> str:
    if ch.isalnum() or ch in " _-:/\"'#@=,;<>!~&%" or ord(ch) > 127:
        if ch.isalpha() and ch.isascii() and rnd.random() < 0.04:
            return "\\x%02x" % ord(ch)
        return ch
    if ch == "\t":
        return "\\t"
    return "\\" + ch


class Node:
    def regex(self) -> str: ...
    def sample(self, rnd: random.Random) -> str: ...


class Lit(Node):
    def __init__(self, text: str, src: str):
        self.text, self.src = text, src

    def regex(self):
        return self.src

    def sample(self, rnd):
        return self.text


class CharSet(Node):
    """A class or class escape: `src` and the characters it can produce in samples."""

    def __init__(self, src: str, members: str):
        self.src, self.members = src, members

    def regex(self):
        return self.src

    def sample(self, rnd):
        return rnd.choice(self.members)


class Cat(Node):
    def __init__(self, kids):
        self.kids = kids

    def regex(self):
        return "".join(k.regex() for k in self.kids)

    def sample(self, rnd):
        return "".join(k.sample(rnd) for k in self.kids)


class Alt(Node):
    def __init__(self, kids, kind):
        self.kids, self.kind = kids, kind

    def regex(self):
        body = "|".join(k.regex() for k in self.kids)
        return {"cap": "(%s)", "non": "(?:%s)", "named": "(?<g>%s)"}[self.kind] % body

    def sample(self, rnd):
        return rnd.choice(self.kids).sample(rnd)


class Flagged(Node):
    """(?i:X) / (?s:X): the sample may flip ASCII case under i."""

    def __init__(self, kid, flags):
        self.kid, self.flags = kid, flags

    def regex(self):
        return "(?%s:%s)" % (self.flags, self.kid.regex())

    def sample(self, rnd):
        s = self.kid.sample(rnd)
        if "i" in self.flags:
            s = "".join(c.swapcase() if c.isascii() and rnd.random() < 0.5 else c for c in s)
        return s


class Rep(Node):
    def __init__(self, kid, lo, hi, suffix):
        self.kid, self.lo, self.hi, self.suffix = kid, lo, hi, suffix

    def regex(self):
        r = self.kid.regex()
        if not isinstance(self.kid, (CharSet, Alt, Flagged)) and not (isinstance(self.kid, Lit) and len(self.kid.text) == 1):
            r = "(?:%s)" % r
        return r + self.suffix

    def sample(self, rnd):
        hi = self.hi if self.hi is not None else self.lo + 3
        return "".join(self.kid.sample(rnd) for _ in range(rnd.randint(self.lo, min(hi, self.lo + 4))))


class QuotedRep(Node):
    """\\Qtext\\E followed by a quantifier: Java binds it to the LAST quoted character."""

    def __init__(self, text, lo, hi, suffix):
        self.text, self.lo, self.hi, self.suffix = text, lo, hi, suffix

    def regex(self):
        return "\\Q" + self.text + "\\E" + self.suffix

    def sample(self, rnd):
        hi = self.hi if self.hi is not None else self.lo + 3
        return self.text[:-1] + self.text[-1] * rnd.randint(self.lo, min(hi, self.lo + 4))


class Assert(Node):
    def __init__(self, src):
        self.src = src

    def regex(self):
        return self.src

    def sample(self, rnd):
        return ""


CLASSES: List[Tuple[str, str]] = [
    (r"\d", "0123456789"), (r"\w", "abzAZ09_"), (r"\s", " \t"), (r"\D", "ab -."), (r"\W", " -.:é"), (r"\S", "ab0-.é"),
    ("[abc]", "abc"), ("[a-cx]", "abcx"), ("[^abc]", "xyz0 -"), ("[0-9a-f]", "09af"), (r"[\w.-]", "az09_.-"), (r"[^\s]", "ab0-"),
    (r"[\d:]", "09:"), (r"\p{Alpha}", "azAZ"), (r"\p{Upper}", "ABZ"), (r"\P{Digit}", "ab -"), (r"\p{Punct}", "-._:/"),
    (".", "ab0 -_é€"), (r"[\]\[]", "[]"), ("[é€]", "é€"), ("[^\\x00-\\x7f]", "é€"), (r"[\Q^-\E]", "^-"),
]


def _literal(rnd: random.Random) -> Node:
    if rnd.random() < 0.5:
        text = rnd.choice(WORDS)
    else:
        text = "".join(rnd.choice(LETTERS + "019 -_.:") for _ in range(rnd.randint(1, 6)))
    if rnd.random() < 0.08:
        return Lit(text, "\\Q" + text + "\\E")
    return Lit(text, "".join(_esc_lit(c, rnd) for c in text))


def _atom(rnd: random.Random, depth: int, in_rep: bool = False) -> Node:
    r = rnd.random()
    if r < 0.45 or depth >= 3:
        return _literal(rnd)
    if r < 0.70:
        return CharSet(*rnd.choice(CLASSES))
    if r < 0.90:
        kids = [_cat(rnd, depth + 1, rnd.randint(1, 2), in_rep) for _ in range(rnd.randint(1, 3))]
        return Alt(kids, rnd.choice(["cap", "non", "non", "named"]))
    return Flagged(_cat(rnd, depth + 1, rnd.randint(1, 2), in_rep), rnd.choice(["i", "i", "s", "is"]))


def _quantified(rnd: random.Random, depth: int, in_rep: bool = False) -> Node:
    # No unbounded quantifier inside a quantified group: (x*)* and friends take a backtracking engine (two of the
    # three engines under test) exponential time on a near miss; the DFA compiler would not mind.
    if rnd.random() < 0.3:
        choices = [(0, 1, "?"), (2, 2, "{2}"), (1, 3, "{1,3}"), (0, 2, "{0,2}")]
        if not in_rep:
            choices += [(0, None, "*"), (1, None, "+"), (2, None, "{2,}")]
        lo, hi, sfx = rnd.choice(choices)
        a = _atom(rnd, depth, True)
        if rnd.random() < 0.15:
            sfx += "?"
        if isinstance(a, Lit) and len(a.text) != 1 and a.src.startswith("\\Q"):
            return QuotedRep(a.text, lo, hi, sfx)
        return Rep(a, lo, hi, sfx)
    return _atom(rnd, depth, in_rep)


def _cat(rnd: random.Random, depth: int, n: int, in_rep: bool = False) -> Node:
    kids: List[Node] = []
    for i in range(n):
        kids.append(_quantified(rnd, depth, in_rep))
        if rnd.random() < 0.08:
            kids.append(Assert(rnd.choice([r"\b", r"\B"])))
    return Cat(kids)


def random_regex(rnd: random.Random) -> Node:
    body = _cat(rnd, 0, rnd.randint(1, 4))
    kids: List[Node] = [body]
    r = rnd.random()
    if r < 0.15:
        kids.insert(0, Assert(rnd.choice(["^", r"\A"])))
    r = rnd.random()
    if r < 0.15:
        kids.append(Assert(rnd.choice(["$", r"\Z", r"\z"])))
    if rnd.random() < 0.1:
        kids.insert(0, Assert("(?i)"))   # sample keeps the case: still a match
    return Cat(kids)


def _mutate(s: str, rnd: random.Random) -> str:
    if not s:
        return rnd.choice(NOISE)
    i = rnd.randrange(len(s))
    op = rnd.random()
    if op < 0.4:
        return s[:i] + s[i + 1:]
    if op < 0.8:
        return s[:i] + rnd.choice(NOISE) + s[i + 1:]
    return s[:i] + rnd.choice(NOISE) + s[i:]


def lines_for(node: Node, rnd: random.Random, n: int) -> List[str]:
    """n lines for one regex: exact samples, embedded samples, mutations, noise.  No '\\n', no
    trailing '\\r' (a line of a log cannot hold them)."""
    out = []
    for k in range(n):
        s = node.sample(rnd)
        kind = k % 4
        if kind == 1:
            s = "".join(rnd.choice(NOISE) for _ in range(rnd.randint(0, 6))) + s + "".join(rnd.choice(NOISE) for _ in range(rnd.randint(0, 6)))
        elif kind == 2:
            s = _mutate(s, rnd)
        elif kind == 3:
            s = "".join(rnd.choice(NOISE) for _ in range(rnd.randint(0, 24)))
        s = s.replace("\n", " ")
        while s.endswith("\r"):
            s = s[:-1]
        out.append(s)
    return out
