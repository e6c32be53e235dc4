"""Random transcript generator for differential tests (adversarial, small)."""
import random

WORDS = ["the", "the", "a", "I", "you", "know", "like", "yeah", "so", "it's", "don't", "we're", "that", "that's",
         "website.com", "e.g.this", "3.5", "1,000.", "Mr.", "J.", "Bach", "ok!!yes", "wait...what", "no-no", "café",
         "你好", "x́", "_a", "U.S.A.", "(hello)", "\"No.\"", "Then", "what?", "really?!", "hmm,", "well;",
         "and:", "100%", "$5", "ﬁne", "Ünï", "straße", "'s", "'LL", "[00:05]", "--", ". . .", "naïve", "٣٤",
         "<|endo", "end.", "Stop!", "go", "Z", "b.", "c.d", "é."]
SPACES = [" ", " ", " ", " ", "  ", "\t", "\n", " ", " ", "　", "\x1c", " \n "]


def rand_text(rng, nwords):
    out = []
    for i in range(nwords):
        w = rng.choice(WORDS)
        if rng.random() < 0.15 and out:
            w = out[-2] if len(out) >= 2 and rng.random() < 0.7 else w   # duplicate previous word
        out.append(w)
        out.append(rng.choice(SPACES) if rng.random() < 0.3 else " ")
    if rng.random() < 0.2:
        out.insert(0, rng.choice(SPACES))
    return "".join(out)


def rand_transcript(rng, nseg=None, long_p=0.05):
    nseg = nseg if nseg is not None else rng.randint(0, 60)
    segs = []
    t = rng.choice([0.0, 0, 12.5, 3590.0, 359990.0])
    nspk = rng.randint(1, 4)
    spk = "S0"
    for i in range(nseg):
        if rng.random() < 0.2:
            spk = rng.choice(["S%d" % k for k in range(nspk)] + ["", "Dr. Who", "é"])
        d = rng.choice([0.5, 1, 2.25, 7.5, 30.0, 61, 130.0, 0.1, 0.07])
        r = rng.random()
        if r < 0.05:
            text = rng.choice(["", "   ", "\n\t", "　"])
        elif r < 0.05 + long_p:
            text = rand_text(rng, rng.randint(40, 300))
        else:
            text = rand_text(rng, rng.randint(1, 12))
        seg = {"start": t, "end": t + d, "text": text, "speaker": spk}
        if rng.random() < 0.02:
            del seg["speaker"]
        segs.append(seg)
        t = t + d + rng.choice([0, 0, 0.25, -0.1])
    return segs
