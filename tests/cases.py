"""Hand-written regex / line cases shared by the CPU and GPU parity tests."""

# every line is checked against every regex of the whole table
REGEXES = [
    r"OutOfMemoryError",
    r"(?i)out of memory",
    r"Timeout after \d+ms waiting for \w+",
    r"^\s+at\s",
    r"(ERROR|FATAL).*database",
    r"[Cc]onnection (refused|reset)",
    r"\bOOM\b",
    r"a$",
    r"foo\z",
    r"bar\Z",
    r"\Qa.b\E+",
    r"\p{Alpha}+\d{2,4}",
    "café",
    r"[^\x00-\x7f]+",
    r"x(?i)y(?-i)z",
    r"\d+\.\d+\.\d+\.\d+:\d+ unreachable",
    r"^ERROR",
    r"exit(ed)? with (code|status) [1-9]\d*",
    r"\[\w+-\d+\] panic: .+",
    r"(?i:Killed) process \d+",
    r"\BNPE\B",
    r"back-?off \d{1,3}(\.\d+)?s",
    r"(?s)a.b",
    r"a.b",
    r"[\w.-]+@[\w-]+\.(com|org)",
    "€\\d+",
    r"\x{1F600}+",
    r"(?<code>E\d{4}) fatal",
    r"^(?:\s+at\s|Caused by:|\s+\.\.\.\s\d+\s(?:more|common frames omitted)|\s+Suppressed:)",
    r"x{3,5}y",
    r"(ab|a)(bc|c)?d",
    r"java\.lang\.\w+Exception",
    r"[\]\[]{2}",
    r"\t\S+\(.*:\d+\)",
    r"segfault at [0-9a-f]+ ip [0-9a-f]+",
    r"été \w+",
    r"(?i)[a-c]{3}-\d",
    r"\P{Alpha}\p{Upper}{3}\P{Alpha}",
    # \Q..\E is rewritten in place by Java: a quantifier after \E binds to the last quoted character
    r"^\Qab\E?c$",
    r"x\Qa.\E{2}y",
    r"[\Q^-\E]+\d",
]

TAB = chr(9)
CR = chr(13)

LINES = [
    "java.lang.OutOfMemoryError: Java heap space", "Out Of MEMORY!", "Timeout after 30ms waiting for lock",
    TAB + "at foo.Bar.baz(Bar.java:12)", "at foo", "   at x", "ERROR: the database died", "FATAL database", "INFO database ERROR",
    "connection reset", "Connection refused", "connection refuse", "an OOM here", "ZOOM", "OOM", "banana", "a" + CR, "a" + CR + CR, "xfoo",
    "foox", "foo", "bar" + CR, "bar ", "bar x", "a.ba.b", "a.b", "axb", "ab12", "ab1", "un café", "cafe", "€", "abc",
    "xYz", "xyz", "xYZ", "10.0.0.1:8080 unreachable", "10.0.0:8080 unreachable", "", " ", "ERROR", " ERROR", "exited with code 137",
    "exit with status 0", "exit with status 10", "[worker-7] panic: runtime error", "[worker-7] panic: ", "killed process 4242",
    "KILLED process x", "xNPEx", "NPE", " NPE ", "back-off 10s", "backoff 1.5s", "backoff 1000s", "a" + CR + "b", "a b", "ab", "ab",
    "mail bob.smith@ex-ample.org now", "bob@ex.net", "€100", "\U0001F600\U0001F600", "\U0001F601", "E1234 fatal",
    "E123 fatal", "Caused by: java.io.IOException", TAB + "... 12 more", TAB + "... 3 common frames omitted", "  Suppressed: x", "Suppressed: x",
    "xxy", "xxxy", "xxxxxxy", "abcd", "acd", "abd", "ad", "abbcd", "java.lang.NullPointerException", "java.lang.Exception",
    "[]", "][", "[x]", TAB + "com.acme.Foo(Foo.java:10)", "segfault at 7f3a ip 00000000004005d4 sp", "segfault at zz ip 1",
    "ac", "c", "abc", "xa..y", "xa.a.y", "xa.y", "^-^5", "a5",
    "x" * 300 + " OutOfMemoryError " + "y" * 5000, "é" * 700 + "connection reset", "été chaud", "ete chaud",
    "AbC-7", "abd-7", " OOM ", "xOOMx", "a ", "foo ", "bar",
]

REJECTED = [
    # (regex, status, byte offset or None)
    ("", -3, 0), ("a*", -3, 0), (r"(a)\1", -3, 3), ("(?=a)b", -3, 0), ("(?!a)b", -3, 0), ("(?<=a)b", -3, 0), ("(?<!a)b", -3, 0),
    ("a++", -3, 2), ("a*+b", -3, 2), ("a?+b", -3, 2), ("a{2}+b", -3, 4), ("(?m)^a", -3, 0), ("(?x)a b", -3, 0), ("(?u)a", -3, 0),
    ("(?d)a", -3, 0), ("(?U)a", -3, 0), ("[a&&b]", -3, 2), ("[a[b]]", -3, 2), (r"\h", -3, 0), (r"\R", -3, 0), (r"\X", -3, 0),
    (r"\G", -3, 0), (r"\v", -3, 0), (r"\N{DIGIT ONE}", -3, 0), (r"\cA", -3, 0), (r"\07", -3, 0), (r"\k<x>", -3, 0), ("a{", -3, 1),
    ("a{,3}", -3, 1), ("a{x}", -3, 1), ("{", -3, 0), ("a$b", -3, 1), ("a$" + chr(92) + "r", -3, 1), ("(a$)+", -3, 2), ("^$", -3, 0), ("(?>a)", -3, 0),
    (r"\p{L}", -3, 0), (r"\p{IsLatin}", -3, 0), (r"\pL", -3, 0), ("a{300}", -4, 1), ("a{2,1000}", -4, 1), ("*a", -2, 0), ("+", -2, 0),
    ("(a", -2, 0), ("a)", -2, 1), ("[a", -2, 0), ("[]a]", -3, 0), ("a**", -2, 2), ("a{3,1}", -2, 1), (r"\q", -2, 0), (chr(92), -2, 0),
    ("[b-a]", -2, 3), (r"\x1", -2, 0), (r"\uD800", -3, 0), (r"^*a", -3, 0), (r"\b+a", -3, 0), ("()", -3, 0), ("a|", -3, 0),
    (r"[\b]", -3, 1), ("x?y?", -3, 0),
]
