"""Randomised parity soak (not collected by pytest): random libraries, pod counts, sizes, CRLF, cut-off
logs and pipeline depths, every result compared with the oracle.  Run on a GPU box:

    python tests/soak_parity.py [seed] [seconds]

Round 1: 757 synthetic configurations / 62 705 events in 200 s and 107 hand-written-regex configurations /
20 291 events in 150 s (both tiers, secondaries, long lines, CRLF), all bit-exact."""
import sys, random, signal, time
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from operator_b200 import Engine, synth
from tests.test_gpu_parity import check_parity

def soak_synthetic(seed, seconds):
    rnd = random.Random(seed)
    eng = Engine(1)
    t0 = time.time(); n = 0; ev = 0
    while time.time() - t0 < seconds:
        npat = rnd.choice([5, 20, 50, 120, 300])
        lib, wit, swit = synth.make_library(npat, seed=rnd.randrange(1 << 30))
        logs = []
        for _ in range(rnd.choice([1, 1, 2, 5, 17])):
            size = rnd.choice([0, 1, 300, 5000, 60000, 250000])
            if size < 300:
                logs.append(b"x" * size)
                continue
            c, _ = synth.make_corpus(size, wit, swit, p_hit=rnd.choice([0, 1e-3, 2e-2, 0.2]), seed=rnd.randrange(1 << 30), crlf=rnd.random() < 0.3)
            raw = bytes(c)
            cut = rnd.choice(["none", "tail", "mid", "nl"])
            if cut == "tail": raw = raw.rstrip(b"\r\n")
            elif cut == "mid" and len(raw) > 2: raw = raw[:rnd.randrange(1, len(raw))]
            elif cut == "nl": raw = raw + b"\n\n\r\n"
            logs.append(raw)
        depth = rnd.choice([1, 1, 3])
        eng.set_pipeline(depth, 0)
        res = check_parity(eng, lib, logs)
        n += 1; ev += res.n_events
    print("soak ok: %d configurations, %d events" % (n, ev))


def soak_handwritten(seed, seconds):
    """Second mode (`python tests/soak_parity.py <seed> <seconds> hand`): random subsets of the hand-written
    regexes (both tiers, secondaries, windows) over logs shuffled together from the hand-written lines,
    random filler and long lines."""
    from tests import cases
    from tests.test_gpu_parity import check_parity, lib_of
    rnd = random.Random(seed)
    eng = Engine(1)
    t0 = time.time(); n = 0; ev = 0
    words = [w for l in cases.LINES for w in l.split()] + ["", " ", "\t", "at", "Caused", "by:", "é", "€"]
    while time.time() - t0 < seconds:
        regs = rnd.sample(cases.REGEXES, rnd.randint(1, min(12, len(cases.REGEXES))))
        lib = lib_of(regs, context_extraction={"lines_before": rnd.randint(0, 6), "lines_after": rnd.randint(0, 8), "include_stack_trace": rnd.random() < 0.5})
        for p in lib["patterns"]:
            if rnd.random() < 0.4:
                p["secondary_patterns"] = [{"regex": rnd.choice(cases.REGEXES), "weight": rnd.choice([0.1, 0.25, 0.5]), "proximity_window": rnd.randint(1, 12)}
                                           for _ in range(rnd.randint(1, 3))]
        logs = []
        for _ in range(rnd.choice([1, 2, 6])):
            lines = []
            for _ in range(rnd.choice([1, 8, 60, 400])):
                r = rnd.random()
                if r < 0.45:
                    lines.append(rnd.choice(cases.LINES))
                elif r < 0.9:
                    lines.append(" ".join(rnd.choice(words) for _ in range(rnd.randint(0, 14))))
                else:
                    lines.append(" ".join(rnd.choice(words) for _ in range(rnd.randint(60, 900))) + " " + rnd.choice(cases.LINES))
            sep = "\r\n" if rnd.random() < 0.3 else "\n"
            logs.append((sep.join(lines) + rnd.choice(["", sep, sep + sep])).encode("utf-8"))
        eng.set_pipeline(rnd.choice([1, 1, 3]), 0)
        try:
            res = check_parity(eng, lib, logs)
        except Exception as e:
            if "nullable" in str(e) or "UNSUPPORTED" in str(e) or "SYNTAX" in str(e):
                continue          # a sampled secondary the compiler rejects (documented subset)
            raise
        n += 1; ev += res.n_events
    print("soak (hand-written) ok: %d configurations, %d events" % (n, ev))


def soak_fuzz(seed, seconds):
    """Third mode (`python tests/soak_parity.py <seed> <seconds> fuzz`): libraries drawn from the grammar-based generator
    of tests/regex_fuzz.py (what the CPU three-way test runs through Python re, PCRE2 and the host DFA walk), as
    primaries and secondaries, over logs made of their own samples, near misses and noise - through the kernels."""
    from tests import regex_fuzz
    from tests.test_gpu_parity import check_parity, lib_of
    rnd = random.Random(seed)
    eng = Engine(1)
    t0 = time.time(); n = 0; ev = 0; skipped = 0; slow = 0

    class OracleTooSlow(Exception):
        pass

    def on_alarm(signum, frame):
        raise OracleTooSlow()

    signal.signal(signal.SIGALRM, on_alarm)
    while time.time() - t0 < seconds:
        nodes = [regex_fuzz.random_regex(rnd) for _ in range(rnd.randint(1, 10))]
        regs = [nd.regex() for nd in nodes]
        lib = lib_of(regs, context_extraction={"lines_before": rnd.randint(0, 4), "lines_after": rnd.randint(0, 6), "include_stack_trace": rnd.random() < 0.3})
        for p in lib["patterns"]:
            if rnd.random() < 0.3:
                p["secondary_patterns"] = [{"regex": regex_fuzz.random_regex(rnd).regex(), "weight": 0.25, "proximity_window": rnd.randint(1, 8)}]
        logs = []
        for _ in range(rnd.choice([1, 3])):
            lines = []
            for nd in nodes:
                lines += regex_fuzz.lines_for(nd, rnd, rnd.choice([4, 16]))
            rnd.shuffle(lines)
            sep = "\r\n" if rnd.random() < 0.3 else "\n"
            logs.append((sep.join(lines) + sep).encode("utf-8"))
        eng.set_pipeline(rnd.choice([1, 1, 3]), 0)
        # The oracle is a backtracking engine (Python re): a drawn regex with nested alternations can take it minutes on a
        # near miss.  The kernels do not care, but there is nothing to compare with then: give the oracle 20 s per configuration.
        signal.alarm(20)
        try:
            res = check_parity(eng, lib, logs)
        except OracleTooSlow:
            slow += 1
            continue
        except Exception as e:
            if any(k in str(e) for k in ("nullable", "UNSUPPORTED", "SYNTAX", "TOO_COMPLEX", "never match")):
                skipped += 1
                continue          # a drawn regex outside the subset (e.g. one that can match the empty string): rejected, as documented
            raise
        finally:
            signal.alarm(0)
        n += 1; ev += res.n_events
    print("soak (fuzz) ok: %d configurations, %d events, %d libraries rejected by the compiler, %d dropped because the oracle was too slow" % (n, ev, skipped, slow))


if __name__ == "__main__":
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    seconds = float(sys.argv[2]) if len(sys.argv) > 2 else 120
    mode = sys.argv[3] if len(sys.argv) > 3 else "synthetic"
    {"hand": soak_handwritten, "fuzz": soak_fuzz}.get(mode, soak_synthetic)(seed, seconds)
