"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

Parser for the Java-regex subset fixed in SPEC.md section 3, plus emitters that restate a parsed
pattern in Python-``re`` and PCRE2 syntax so that two independent backtracking engines can
decide ``Pattern.compile(rx).matcher(line).find()`` for the oracle.

Parity unpinned: the reference tree holds no regex code and no tests;
the only in-tree anchors are the call sites LogParserRestClient.java:37-39 and
LogParserClient.java:40-41. java.util.regex (JDK 21, the runtime named in pom.xml:10) is the
third-party engine whose published semantics this file follows.

This is a second, independent implementation of the grammar that
operator_b200/csrc/pm_regex.cc implements in C++; tests diff their accept/reject verdicts.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Tuple

MAX_CP = 0x10FFFF
MAX_REPEAT = 256

E_UNSUPPORTED = "unsupported"
E_SYNTAX = "syntax"
E_TOO_COMPLEX = "too_complex"


class JavaRegexError(Exception):
    def __init__(self, kind: str, offset: int, msg: str):
        super().__init__(f"{kind} at {offset}: {msg}")
        self.kind, self.offset, self.msg = kind, offset, msg


# --------------------------------------------------------------------------- AST
@dataclass
class Node:
    kind: str                      # lit set dot cat alt rep bol eolj end wb nwb empty
    cp: int = 0
    ci: bool = False
    ranges: List[Tuple[int, int]] = field(default_factory=list)   # for 'set' (already final)
    dotall: bool = False
    kids: List["Node"] = field(default_factory=list)
    lo: int = 0
    hi: Optional[int] = None
    lazy: bool = False
    pos: int = 0


def _norm(ranges):
    out = []
    for lo, hi in sorted(ranges):
        if out and lo <= out[-1][1] + 1:
            out[-1] = (out[-1][0], max(out[-1][1], hi))
        else:
            out.append((lo, hi))
    return out


def _negate(ranges):
    out, prev = [], 0
    for lo, hi in _norm(ranges):
        if lo > prev:
            out.append((prev, lo - 1))
        prev = hi + 1
    if prev <= MAX_CP:
        out.append((prev, MAX_CP))
    return out


def _casefold_ascii(ranges):
    extra = []
    for lo, hi in ranges:
        a, b = max(lo, 0x41), min(hi, 0x5A)
        if a <= b:
            extra.append((a + 32, b + 32))
        a, b = max(lo, 0x61), min(hi, 0x7A)
        if a <= b:
            extra.append((a - 32, b - 32))
    return _norm(list(ranges) + extra)


DIGIT = [(0x30, 0x39)]
WORD = [(0x30, 0x39), (0x41, 0x5A), (0x5F, 0x5F), (0x61, 0x7A)]
SPACE = [(0x09, 0x0D), (0x20, 0x20)]
POSIX = {
    "Alpha": [(0x41, 0x5A), (0x61, 0x7A)],
    "Digit": DIGIT,
    "Alnum": [(0x30, 0x39), (0x41, 0x5A), (0x61, 0x7A)],
    "Upper": [(0x41, 0x5A)],
    "Lower": [(0x61, 0x7A)],
    "Punct": [(0x21, 0x2F), (0x3A, 0x40), (0x5B, 0x60), (0x7B, 0x7E)],
    "Space": SPACE,
    "XDigit": [(0x30, 0x39), (0x41, 0x46), (0x61, 0x66)],
}
LINE_TERM = [(0x0A, 0x0A), (0x0D, 0x0D), (0x85, 0x85), (0x2028, 0x2029)]
SIMPLE_ESC = {"t": 9, "n": 10, "r": 13, "f": 12, "a": 7, "e": 27}
REJECT_ESC = set("GRXhHvVNck0123456789")


@dataclass
class _Flags:
    ci: bool = False
    dotall: bool = False


def _remove_qe_quoting(src: str):
    """java.util.regex.Pattern rewrites the pattern text before it parses it (Pattern.RemoveQEQuoting
    in the JDK sources; behaviour documented under "Quotation" in the Pattern Javadoc): from the first
    \\Q on, quoted characters become individually escaped characters IN PLACE.  Consequences the
    parser must reproduce: a quantifier after \\E binds to the LAST quoted character only
    (`\\Qab\\E?c` is `ab?c`), quoting works inside a character class, ASCII letters are copied as they
    are, a digit that opens a quotation becomes \\x3N (so it cannot extend an escape in front of it),
    every other ASCII character gets a backslash.  Returns the rewritten text and, per character of
    it, the index of the source character it came from (error offsets refer to the source)."""
    n = len(src)
    i = 0
    while i < n - 1:
        if src[i] != "\\":
            i += 1
        elif src[i + 1] != "Q":
            i += 2
        else:
            break
    if i >= n - 1:
        return src, list(range(n))
    out = list(src[:i])
    origin = list(range(i))
    i += 2
    in_quote, begin = True, True

    def put(ch, at):
        out.append(ch)
        origin.append(at)

    while i < n:
        at = i
        c = src[i]
        i += 1
        if ord(c) >= 128 or (c.isalpha() and c.isascii()):
            put(c, at)
        elif c.isdigit() and c.isascii():
            if begin:
                for ch in "\\x3":
                    put(ch, at)
            put(c, at)
        elif c != "\\":
            if in_quote:
                put("\\", at)
            put(c, at)
        elif in_quote:
            if i < n and src[i] == "E":
                i += 1
                in_quote = False
            else:
                put("\\", at)
                put("\\", at)
        else:
            if i < n and src[i] == "Q":
                i += 1
                in_quote, begin = True, True
                continue
            put(c, at)
            if i < n:
                put(src[i], i)
                i += 1
        begin = False
    return "".join(out), origin


class _Parser:
    def __init__(self, src: str):
        self.src_len = len(src)
        self.s, self.origin = _remove_qe_quoting(src)
        self.i = 0

    # -- helpers
    def _peek(self, k=0):
        j = self.i + k
        return self.s[j] if j < len(self.s) else ""

    def _err(self, kind, msg, at=None):
        at = self.i if at is None else at
        raise JavaRegexError(kind, self.origin[at] if at < len(self.origin) else self.src_len, msg)

    # -- grammar
    def parse(self) -> Node:
        if not self.s:
            self._err(E_UNSUPPORTED, "empty regex")
        fl = _Flags()
        n = self._alt(fl, top=True)
        if self.i < len(self.s):
            self._err(E_SYNTAX, "unmatched ')'")
        return n

    def _alt(self, fl: _Flags, top=False) -> Node:
        start = self.i
        alts = [self._cat(fl)]
        while self._peek() == "|":
            self.i += 1
            alts.append(self._cat(fl))
        return alts[0] if len(alts) == 1 else Node("alt", kids=alts, pos=start)

    def _cat(self, fl: _Flags) -> Node:
        start = self.i
        items: List[Node] = []
        while self.i < len(self.s) and self._peek() not in "|)":
            n = self._repeat(fl)
            if n is not None:
                items.append(n)
        if not items:
            return Node("empty", pos=start)
        return items[0] if len(items) == 1 else Node("cat", kids=items, pos=start)

    def _repeat(self, fl: _Flags) -> Optional[Node]:
        at = self.i
        atom = self._atom(fl)
        if atom is None:          # inline flag group, produced nothing
            return None
        while True:
            c = self._peek()
            if c == "*":
                lo, hi = 0, None
                self.i += 1
            elif c == "+":
                lo, hi = 1, None
                self.i += 1
            elif c == "?":
                lo, hi = 0, 1
                self.i += 1
            elif c == "{":
                lo, hi = self._braces()
            else:
                return atom
            if atom.kind in ("bol", "eolj", "end", "wb", "nwb"):
                self._err(E_UNSUPPORTED, "quantified assertion", at)
            lazy = False
            if self._peek() == "?":
                lazy = True
                self.i += 1
            elif self._peek() == "+":
                self._err(E_UNSUPPORTED, "possessive quantifier")
            if self._peek() in ("*", "+", "?", "{") and self._peek() != "":
                # Java: "Dangling meta character"
                self._err(E_SYNTAX, "dangling quantifier")
            atom = Node("rep", kids=[atom], lo=lo, hi=hi, lazy=lazy, pos=at)
            return atom

    def _braces(self):
        at = self.i
        self.i += 1
        j = self.i
        while self._peek().isdigit() and self._peek().isascii():
            self.i += 1
        if j == self.i:
            self._err(E_UNSUPPORTED, "'{' is not a valid quantifier", at)
        lo = int(self.s[j:self.i])
        hi: Optional[int] = lo
        if self._peek() == ",":
            self.i += 1
            j = self.i
            while self._peek().isdigit() and self._peek().isascii():
                self.i += 1
            hi = int(self.s[j:self.i]) if self.i > j else None
        if self._peek() != "}":
            self._err(E_UNSUPPORTED, "'{' is not a valid quantifier", at)
        self.i += 1
        if hi is not None and hi < lo:
            self._err(E_SYNTAX, "repetition bounds out of order", at)
        if lo > MAX_REPEAT or (hi is not None and hi > MAX_REPEAT):
            self._err(E_TOO_COMPLEX, "repetition bound > %d" % MAX_REPEAT, at)
        return lo, hi

    def _lit(self, cp: int, fl: _Flags, at: int) -> Node:
        if 0xD800 <= cp <= 0xDFFF:
            self._err(E_UNSUPPORTED, "surrogate code point", at)
        ci = fl.ci and (0x41 <= cp <= 0x5A or 0x61 <= cp <= 0x7A)
        return Node("lit", cp=cp, ci=ci, pos=at)

    def _set(self, ranges, fl: _Flags, at: int, negate=False) -> Node:
        r = _norm(ranges)
        if fl.ci:
            r = _casefold_ascii(r)
        if negate:
            r = _negate(r)
        return Node("set", ranges=r, pos=at)

    def _atom(self, fl: _Flags) -> Optional[Node]:
        at = self.i
        c = self._peek()
        if c == "(":
            return self._group(fl)
        if c == "[":
            return self._klass(fl)
        if c == ".":
            self.i += 1
            return Node("dot", dotall=fl.dotall, pos=at)
        if c == "^":
            self.i += 1
            return Node("bol", pos=at)
        if c == "$":
            self.i += 1
            return Node("eolj", pos=at)
        if c in "*+?":
            self._err(E_SYNTAX, "dangling quantifier")
        if c == "{":
            self._err(E_UNSUPPORTED, "'{' is not a valid quantifier")
        if c == "\\":
            return self._escape(fl)
        self.i += 1
        return self._lit(ord(c), fl, at)

    def _group(self, fl: _Flags) -> Optional[Node]:
        at = self.i
        self.i += 1
        inner = _Flags(fl.ci, fl.dotall)
        if self._peek() == "?":
            self.i += 1
            c = self._peek()
            if c == ":":
                self.i += 1
            elif c == "<":
                n1 = self._peek(1)
                if n1 in ("=", "!"):
                    self._err(E_UNSUPPORTED, "look-behind", at)
                self.i += 1
                j = self.i
                while self._peek().isascii() and self._peek().isalnum():
                    self.i += 1
                if j == self.i or not self.s[j].isalpha() or self._peek() != ">":
                    self._err(E_SYNTAX, "bad group name", at)
                self.i += 1
            elif c in ("=", "!"):
                self._err(E_UNSUPPORTED, "look-ahead", at)
            elif c == ">":
                self._err(E_UNSUPPORTED, "atomic group", at)
            else:
                on = True
                seen = False
                tmp = _Flags(fl.ci, fl.dotall)
                while True:
                    c = self._peek()
                    if c == "i":
                        tmp.ci = on
                    elif c == "s":
                        tmp.dotall = on
                    elif c == "-" and on:
                        on = False
                    elif c in (")", ":") and (seen or True):
                        break
                    elif c == "":
                        self._err(E_SYNTAX, "unterminated group", at)
                    else:
                        self._err(E_UNSUPPORTED, "inline flag '%s'" % c, at)
                    seen = True
                    self.i += 1
                if self._peek() == ")":
                    self.i += 1
                    fl.ci, fl.dotall = tmp.ci, tmp.dotall      # rest of the enclosing group
                    return None
                self.i += 1                                     # ':'
                inner = tmp
        n = self._alt(inner)
        if self._peek() != ")":
            self._err(E_SYNTAX, "unterminated group", at)
        self.i += 1
        return Node("cat", kids=[n], pos=at) if n.kind == "empty" else n

    def _hex(self, n, at):
        h = self.s[self.i:self.i + n]
        if len(h) != n or any(ch not in "0123456789abcdefABCDEF" for ch in h):
            self._err(E_SYNTAX, "bad hex escape", at)
        self.i += n
        return int(h, 16)

    def _escape_cp(self, at) -> Optional[int]:
        """After the backslash: escapes that denote ONE code point (shared by class parser)."""
        c = self._peek()
        if c in SIMPLE_ESC:
            self.i += 1
            return SIMPLE_ESC[c]
        if c == "x":
            self.i += 1
            if self._peek() == "{":
                self.i += 1
                j = self.i
                while self._peek() and self._peek() in "0123456789abcdefABCDEF":
                    self.i += 1
                if j == self.i or self._peek() != "}" or self.i - j > 6:
                    self._err(E_SYNTAX, "bad hex escape", at)
                v = int(self.s[j:self.i], 16)
                self.i += 1
                if v > MAX_CP:
                    self._err(E_SYNTAX, "code point out of range", at)
                return v
            return self._hex(2, at)
        if c == "u":
            self.i += 1
            return self._hex(4, at)
        return None

    def _posix(self, at):
        neg = self._peek() == "P"
        self.i += 1
        if self._peek() != "{":
            self._err(E_UNSUPPORTED, "\\p without braces", at)
        j = self.s.find("}", self.i)
        if j < 0:
            self._err(E_SYNTAX, "unterminated \\p{", at)
        name = self.s[self.i + 1:j]
        self.i = j + 1
        if name not in POSIX:
            self._err(E_UNSUPPORTED, "property \\p{%s}" % name, at)
        return POSIX[name], neg

    def _escape(self, fl: _Flags) -> Node:
        at = self.i
        self.i += 1
        c = self._peek()
        if c == "":
            self._err(E_SYNTAX, "trailing backslash", at)
        cp = self._escape_cp(at)
        if cp is not None:
            return self._lit(cp, fl, at)
        if c in "dwsDWS":
            self.i += 1
            base = {"d": DIGIT, "w": WORD, "s": SPACE}[c.lower()]
            # (?i) cannot change these sets: they are closed under ASCII case folding
            return Node("set", ranges=_negate(base) if c.isupper() else _norm(base), pos=at)
        if c in "pP":
            r, neg = self._posix(at)
            return self._set(r, fl, at, negate=neg)
        if c == "b":
            self.i += 1
            if self._peek() == "{":
                self._err(E_UNSUPPORTED, "\\b{..}", at)
            return Node("wb", pos=at)
        if c == "B":
            self.i += 1
            return Node("nwb", pos=at)
        if c == "A":
            self.i += 1
            return Node("bol", pos=at)
        if c == "Z":
            self.i += 1
            return Node("eolj", pos=at)
        if c == "z":
            self.i += 1
            return Node("end", pos=at)
        if c in REJECT_ESC:
            self._err(E_UNSUPPORTED, "escape \\%s" % c, at)
        if c.isascii() and c.isalpha():
            self._err(E_SYNTAX, "illegal escape \\%s" % c, at)
        self.i += 1
        return self._lit(ord(c), fl, at)

    def _klass(self, fl: _Flags) -> Node:
        at = self.i
        self.i += 1
        neg = False
        if self._peek() == "^":
            neg = True
            self.i += 1
        if self._peek() == "]":
            self._err(E_UNSUPPORTED, "']' first in class", at)
        ranges: List[Tuple[int, int]] = []

        def item():
            """Returns ('cp', v) or ('set', ranges)."""
            p = self.i
            c = self._peek()
            if c == "":
                self._err(E_SYNTAX, "unterminated class", at)
            if c == "[":
                self._err(E_UNSUPPORTED, "nested class", p)
            if c == "&" and self._peek(1) == "&":
                self._err(E_UNSUPPORTED, "class intersection", p)
            if c == "\\":
                self.i += 1
                e = self._peek()
                if e == "":
                    self._err(E_SYNTAX, "trailing backslash", p)
                v = self._escape_cp(p)
                if v is not None:
                    return ("cp", v)
                if e in "dwsDWS":
                    self.i += 1
                    base = {"d": DIGIT, "w": WORD, "s": SPACE}[e.lower()]
                    return ("set", _negate(base) if e.isupper() else base)
                if e in "pP":
                    r, n = self._posix(p)
                    return ("set", _negate(r) if n else r)
                if e.isascii() and e.isalnum():
                    self._err(E_UNSUPPORTED, "escape \\%s in class" % e, p)
                self.i += 1
                return ("cp", ord(e))
            self.i += 1
            return ("cp", ord(c))

        while True:
            if self._peek() == "]":
                self.i += 1
                break
            k, v = item()
            if k == "cp" and self._peek() == "-" and self._peek(1) not in ("]", ""):
                self.i += 1
                p2 = self.i
                k2, v2 = item()
                if k2 != "cp":
                    self._err(E_SYNTAX, "bad range end", p2)
                if v2 < v:
                    self._err(E_SYNTAX, "range out of order", p2)
                ranges.append((v, v2))
            elif k == "cp":
                ranges.append((v, v))
            else:
                ranges.extend(v)
        for lo, hi in ranges:
            if lo <= 0xDFFF and hi >= 0xD800 and (lo >= 0xD800 or hi <= 0xDFFF):
                self._err(E_UNSUPPORTED, "surrogate code point", at)
        return self._set(ranges, fl, at, negate=neg)


# --------------------------------------------------------------------------- analysis
INF = 1 << 30


def width(n: Node) -> Tuple[int, int]:
    """(min, max) match width in code points; max INF when unbounded."""
    k = n.kind
    if k in ("lit", "set", "dot"):
        return (1, 1)
    if k in ("bol", "eolj", "end", "wb", "nwb", "empty"):
        return (0, 0)
    if k == "cat":
        lo = hi = 0
        for c in n.kids:
            a, b = width(c)
            lo, hi = lo + a, min(INF, hi + b)
        return (lo, hi)
    if k == "alt":
        ws = [width(c) for c in n.kids]
        return (min(w[0] for w in ws), max(w[1] for w in ws))
    if k == "rep":
        a, b = width(n.kids[0])
        hi = INF if (n.hi is None and b > 0) else min(INF, b * (n.hi or 0))
        return (a * n.lo, hi)
    raise AssertionError(k)


def _check_tail(n: Node, tail: bool):
    """'$' is only supported where nothing consuming can follow it (SPEC section 3)."""
    k = n.kind
    if k == "eolj" and not tail:
        raise JavaRegexError(E_UNSUPPORTED, n.pos, "'$' followed by a consuming item")
    if k == "cat":
        for idx, c in enumerate(n.kids):
            rest_max = sum(width(x)[1] for x in n.kids[idx + 1:])
            _check_tail(c, tail and rest_max == 0)
    elif k == "alt":
        for c in n.kids:
            _check_tail(c, tail)
    elif k == "rep":
        single = n.hi is not None and n.hi <= 1
        _check_tail(n.kids[0], tail and (single or width(n.kids[0])[1] == 0))


def parse(src: str) -> Node:
    n = _Parser(src).parse()
    _check_tail(n, True)
    if width(n)[0] == 0:
        raise JavaRegexError(E_UNSUPPORTED, 0, "regex can match the empty string")
    return n


# --------------------------------------------------------------------------- emitters
def _cp_py(cp: int) -> str:
    return "\\x%02x" % cp if cp < 0x100 else ("\\u%04x" % cp if cp < 0x10000 else "\\U%08x" % cp)


def _cp_pcre(cp: int) -> str:
    return "\\x{%x}" % cp


def _emit(n: Node, cpf, end_tok: str) -> str:
    k = n.kind

    def klass(ranges):
        if not ranges:
            return "(?!)"
        if ranges == [(0, MAX_CP)]:
            return "(?s:.)"
        return "[" + "".join(cpf(a) if a == b else cpf(a) + "-" + cpf(b) for a, b in ranges) + "]"

    if k == "lit":
        if n.ci:
            return klass(_norm([(n.cp, n.cp), (n.cp ^ 0x20, n.cp ^ 0x20)]))
        return cpf(n.cp)
    if k == "set":
        return klass(n.ranges)
    if k == "dot":
        return "(?s:.)" if n.dotall else klass(_negate(LINE_TERM))
    if k == "empty":
        return ""
    if k == "bol":
        return "\\A"
    if k == "end":
        return end_tok
    if k == "eolj":
        return "(?=" + klass(_norm([(0x0D, 0x0D), (0x85, 0x85), (0x2028, 0x2029)])) + "?" + end_tok + ")"
    if k == "wb":
        return "\\b"
    if k == "nwb":
        return "\\B"
    if k == "cat":
        return "".join(_emit(c, cpf, end_tok) if c.kind != "alt" else "(?:" + _emit(c, cpf, end_tok) + ")"
                       for c in n.kids)
    if k == "alt":
        return "|".join(_emit(c, cpf, end_tok) for c in n.kids)
    if k == "rep":
        inner = "(?:" + _emit(n.kids[0], cpf, end_tok) + ")"
        if (n.lo, n.hi) == (0, None):
            q = "*"
        elif (n.lo, n.hi) == (1, None):
            q = "+"
        elif (n.lo, n.hi) == (0, 1):
            q = "?"
        elif n.hi is None:
            q = "{%d,}" % n.lo
        elif n.hi == n.lo:
            q = "{%d}" % n.lo
        else:
            q = "{%d,%d}" % (n.lo, n.hi)
        return inner + q + ("?" if n.lazy else "")
    raise AssertionError(k)


def to_python(src: str) -> str:
    """Java regex (subset) -> equivalent Python ``re`` source; compile with re.ASCII."""
    return _emit(parse(src), _cp_py, "\\Z")


def to_pcre2(src: str) -> str:
    """Java regex (subset) -> equivalent PCRE2 source; compile with PCRE2_UTF only."""
    return _emit(parse(src), _cp_pcre, "\\z")
