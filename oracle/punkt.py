"""Untrained Punkt sentence tokenizer, restated (TEST INFRASTRUCTURE; parity unpinned).

The reference constructs `nltk.tokenize.PunktSentenceTokenizer()` with no training text
(/root/reference/big_chunkeroosky.py:44) and calls `.tokenize(text)` at :334.  nltk==3.8.1
(/root/reference/requirements.txt:1) is not installed and not installable offline, so this module
restates the published algorithm (Kiss & Strunk 2006 as implemented in nltk/tokenize/punkt.py 3.8.1)
for the *empty-parameter* case: no abbreviations, no collocations, no sentence starters, all
orthographic contexts zero.  Nothing in the reference's own tests pins this boundary.

Structure follows the nltk method names so a maintainer can diff it against a real nltk wheel:
  PunktLanguageVars            -> the regexes below
  _match_potential_end_contexts-> potential_end_contexts()
  text_contains_sentbreak      -> context_has_break()
  _annotate_first/second_pass  -> annotate()
  _slices_from_text            -> slices_from_text()
  _realign_boundaries          -> realign()
"""
import re
import string

SENT_END = (".", "?", "!")
PUNCTUATION = (";", ":", ",", ".", "!", "?")
_NONWORD = r"(?:[)\";}\]\*:@\'\({\[\?!])"
_MULTI = r"(?:\-{2,}|\.{2,}|(?:\.\s){2,}\.)"
_WORDSTART = r"[^\(\"\`{\[:;&\#\*@\)}\]\-,]"
WORD_RE = re.compile(
    r"(" + _MULTI + r"|(?=" + _WORDSTART + r")\S+?(?=\s|$|" + _NONWORD + r"|" + _MULTI +
    r"|,(?=$|\s|" + _NONWORD + r"|" + _MULTI + r"))|\S)", re.UNICODE)
PERIOD_CONTEXT_RE = re.compile(
    r"[\.\?!](?=(?P<after_tok>" + _NONWORD + r"|\s+(?P<next_tok>\S+)))", re.UNICODE)
REALIGN_RE = re.compile(r'["\')\]}]+?(?:\s+|(?=--)|$)', re.MULTILINE)
_RE_ELLIPSIS = re.compile(r"\.\.+$")
_RE_NUMERIC = re.compile(r"^-?[\.,]?\d[\d,\.-]*\.?$")
_RE_INITIAL = re.compile(r"[^\W\d]\.$", re.UNICODE)


def word_tokenize(line):
    return WORD_RE.findall(line)


def tokenize_words(text):
    for line in text.split("\n"):
        if line.strip():
            for tok in word_tokenize(line):
                yield tok


def annotate(toks):
    """Return list of sentbreak flags (first pass then second pass) for empty parameters."""
    brk = []
    for tok in toks:
        if tok in SENT_END:
            brk.append(True)
        elif _RE_ELLIPSIS.match(tok):
            brk.append(False)
        elif tok.endswith(".") and not tok.endswith(".."):
            brk.append(True)
        else:
            brk.append(False)
    for i in range(len(toks) - 1):
        t1, t2 = toks[i], toks[i + 1]
        if not t1.endswith("."):
            continue
        is_initial = bool(_RE_INITIAL.match(t1))
        is_number = _RE_NUMERIC.sub("##number##", t1.lower()) == "##number##"
        if is_initial or is_number:
            if t2 in PUNCTUATION or t2[0].islower():
                brk[i] = False          # orthographic heuristic says "not a sentence starter"
            elif is_initial and t2[0].isupper():
                brk[i] = False          # "J. Bach" special case (ortho context never saw lower case)
    return brk


def context_has_break(context):
    toks = list(tokenize_words(context))
    brk = annotate(toks)
    return any(brk[:-1])


def _last_ascii_ws(s):
    for i in range(len(s) - 1, -1, -1):
        if s[i] in string.whitespace:
            return i
    return 0


def potential_end_contexts(text):
    prev_start, prev_stop = 0, 0
    prev_match = None
    for m in PERIOD_CONTEXT_RE.finditer(text):
        before = text[prev_stop:m.start()]
        k = _last_ascii_ws(before)
        if k:
            k += prev_stop + 1
        else:
            k = prev_start
        if prev_match is not None and prev_stop <= k:
            yield prev_match, text[prev_start:prev_stop] + prev_match.group() + prev_match.group("after_tok")
        prev_match = m
        prev_start, prev_stop = k, m.start()
    if prev_match is not None:
        yield prev_match, text[prev_start:prev_stop] + prev_match.group() + prev_match.group("after_tok")


def slices_from_text(text):
    last = 0
    for m, ctx in potential_end_contexts(text):
        if context_has_break(ctx):
            yield (last, m.end())
            last = m.start("next_tok") if m.group("next_tok") else m.end()
    yield (last, len(text.rstrip()))


def realign(text, slices):
    slices = list(slices)
    shift = 0
    for i, (s1a, s1b) in enumerate(slices):
        s1a += shift
        if i + 1 == len(slices):
            if text[s1a:s1b]:
                yield (s1a, s1b)
            continue
        s2a, s2b = slices[i + 1]
        m = REALIGN_RE.match(text[s2a:s2b])
        if m:
            yield (s1a, s2a + len(m.group(0).rstrip()))
            shift = m.end()
        else:
            shift = 0
            if text[s1a:s1b]:
                yield (s1a, s1b)


class PunktSentenceTokenizer:
    def __init__(self, train_text=None, verbose=False, lang_vars=None, token_cls=None):
        if train_text:
            raise NotImplementedError("oracle restates the untrained tokenizer only")

    def span_tokenize(self, text, realign_boundaries=True):
        sl = slices_from_text(text)
        if realign_boundaries:
            sl = realign(text, sl)
        for a, b in sl:
            yield (a, b)

    def tokenize(self, text, realign_boundaries=True):
        return [text[a:b] for a, b in self.span_tokenize(text, realign_boundaries)]
