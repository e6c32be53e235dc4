"""Chunker oracle (TEST INFRASTRUCTURE): restates /root/reference/big_chunkeroosky.py:46-567 as a flat
item stream + two greedy rules, the form the CUDA kernels implement.
Pinned against the real big_chunkeroosky.py (run with the nltk shim and the repo rank table) by
tools/make_golden.py / tests/test_oracle_vs_golden.py.
"""
import re
from .preprocess import format_ts
from . import punkt as _punkt
from . import tokenizer as _tk

_CLAUSE = re.compile(r"([^,.;:?!]+[,.;:?!]+)")


class ChunkerOracle:
    def __init__(self, max_tokens_per_chunk=4000, overlap_tokens=200, tokenizer_name="cl100k_base",
                 context_tokens=150, count=None):
        self.max_tokens_per_chunk = max_tokens_per_chunk
        self.overlap_tokens = overlap_tokens            # stored, never used (big_chunkeroosky.py:40)
        self.context_tokens = context_tokens
        self.effective_max_tokens = max_tokens_per_chunk - context_tokens
        if count is None:
            enc = _tk.encoding()
            count = lambda s: len(enc.encode(s))        # default disallowed_special="all" -> ValueError
        self.count = count
        self.sent = _punkt.PunktSentenceTokenizer()

    # ---- large-segment expansion (big_chunkeroosky.py:267-542) -> list of (segment_dict, text, tokens)
    def _greedy_pieces(self, pieces, mk, E):
        """Shared greedy of :296-329 and :496-540: pieces = [(text, tokens, start, end)]."""
        subs = []
        cur = None
        for text, tok, st, en in pieces:
            if cur is None:
                cur = mk(None)
            if cur["tok"] + tok > E and cur["tok"] > 0:
                subs.append(cur)
                cur = mk((st, en, text, tok))
                if cur["tok"]:
                    continue
            cur["text"] = cur["text"] + " " + text if cur["text"] else text
            cur["tok"] += tok
            cur["seg"]["end"] = en
            cur["seg"]["text"] = cur["text"]
        if cur is not None and cur["tok"] > 0:
            subs.append(cur)
        return subs

    def _expand_combined(self, seg):
        E = self.effective_max_tokens
        pieces = []
        for ts in seg["segment_timestamps"]:
            t = "[" + format_ts(ts["start"]) + "] " + ts["text"]
            pieces.append((t, self.count(t), ts["start"], ts["end"]))
        state = {"first": True}

        def mk(init):
            # :282-294 first sub-chunk starts at the first piece; :305-317 later ones at the piece that
            # overflowed, and begin EMPTY (the piece is then appended by the common code)
            st = seg["segment_timestamps"][0]["start"] if state["first"] else init[0]
            state["first"] = False
            return {"seg": {"start": st, "end": None, "speaker": seg["speaker"], "text": "",
                            "is_sub_chunk": True, "parent_segment_start": seg["start"],
                            "parent_segment_end": seg["end"]}, "text": "", "tok": 0}
        return self._greedy_pieces(pieces, mk, E)

    def _expand_clauses(self, sentence, t_start, t_end):
        E = self.effective_max_tokens
        clauses = _CLAUSE.findall(sentence)
        if not clauses:
            words = sentence.split()
            clauses = [" ".join(words[i:i + 20]) for i in range(0, len(words), 20)]
        span = t_end - t_start
        tpc = span / len(sentence) if len(sentence) > 0 else 0
        done = 0
        pieces = []
        for cl in clauses:
            c = cl.strip()
            if not c:
                continue
            c_start = t_start + (tpc * done)
            c_end = c_start + tpc * len(c)
            done += len(c)
            t = "[" + format_ts(c_start) + "] " + c
            pieces.append((t, self.count(t), c_start, c_end))
        state = {"first": True}

        def mk(init):
            if state["first"]:
                state["first"] = False
                return {"seg": {"start": t_start, "end": None, "speaker": "", "text": "",
                                "is_sub_chunk": True, "is_clause": True}, "text": "", "tok": 0}
            st, en, text, tok = init     # :517-528 the overflowing clause seeds the new sub-chunk
            return {"seg": {"start": st, "end": en, "speaker": "", "text": text,
                            "is_sub_chunk": True, "is_clause": True}, "text": text, "tok": tok}
        return self._greedy_pieces(pieces, mk, E)

    def _expand_sentences(self, seg):
        E = self.effective_max_tokens
        text = seg["text"]
        tpc = (seg["end"] - seg["start"]) / len(text) if len(text) > 0 else 0
        done = 0

        def fresh(st, en=None, tx="", tok=0):
            return {"seg": {"start": st, "end": en, "speaker": seg["speaker"], "text": tx,
                            "is_sub_chunk": True, "parent_segment_start": seg["start"],
                            "parent_segment_end": seg["end"]}, "text": tx, "tok": tok}
        subs = []
        cur = fresh(seg["start"])
        for sentence in self.sent.tokenize(text):
            s = sentence.strip()
            if not s:
                continue
            s_start = seg["start"] + (tpc * done)
            s_end = s_start + tpc * len(s)
            done += len(s)
            f = "[" + format_ts(s_start) + "] " + s
            tok = self.count(f)
            if tok > E:
                cl = self._expand_clauses(s, s_start, s_end)
                if cur["tok"] > 0:
                    cur["seg"]["end"] = s_start
                    cur["seg"]["text"] = cur["text"]
                    subs.append(cur)
                subs.extend(cl)
                cur = fresh(s_end)
            elif cur["tok"] + tok > E and cur["tok"] > 0:
                cur["seg"]["end"] = s_start
                cur["seg"]["text"] = cur["text"]
                subs.append(cur)
                cur = fresh(s_start, s_end, f, tok)
            else:
                cur["text"] = cur["text"] + " " + f if cur["text"] else f
                cur["tok"] += tok
                cur["seg"]["end"] = s_end
                cur["seg"]["text"] = cur["text"]
        if cur["tok"] > 0:
            subs.append(cur)
        return subs

    def expand(self, seg):
        if "segment_timestamps" in seg and seg.get("is_combined", False):
            subs = self._expand_combined(seg)
        else:
            subs = self._expand_sentences(seg)
        return [(s["seg"], s["text"], s["tok"]) for s in subs]

    # ---- packing (big_chunkeroosky.py:46-145) + finalize (:147-232) + postprocess (:544-567)
    def chunk(self, segs, add_context=True):
        if not segs:
            return []
        E = self.effective_max_tokens
        chunks = []

        def new(start, seg=None, text="", tok=0):
            return {"segments": [seg] if seg is not None else [], "text": text, "token_count": tok,
                    "start_time": start, "end_time": seg["end"] if seg is not None else None,
                    "speakers": {seg["speaker"]} if seg is not None else set()}

        def finalize(ch):
            ch["speakers"] = sorted(ch["speakers"])
            ch["chunk_index"] = len(chunks)
            ch["total_chunks"] = None
            first, last = ch["segments"][0]["start"], ch["segments"][-1]["end"]
            t0 = chunks[0]["segments"][0]["start"] if chunks else first
            ch["position_percentage"] = (first - t0) / (last - t0) * 100 if last > t0 else 0
            if add_context:
                hdr = ("--- TRANSCRIPT CHUNK INFORMATION ---\n"
                       "Time Range: " + format_ts(ch["start_time"]) + " - " + format_ts(ch["end_time"]) + "\n"
                       "Speakers: " + ", ".join(ch["speakers"]) + "\n"
                       "Position: Chunk %d (approximately %.1f%% through the transcript)\n"
                       "--- TRANSCRIPT CHUNK CONTENT ---" % (ch["chunk_index"] + 1, ch["position_percentage"]))
                ch["text_with_context"] = hdr + "\n\n" + ch["text"]
            else:
                ch["text_with_context"] = ch["text"]
            chunks.append(ch)

        def append(ch, seg, text, tok):
            ch["segments"].append(seg)
            ch["text"] = ch["text"] + "\n\n" + text if ch["text"] else text
            ch["token_count"] += tok
            ch["end_time"] = seg["end"]
            ch["speakers"].add(seg["speaker"])

        cur = new(segs[0]["start"])
        for seg in segs:
            ftext = "[" + format_ts(seg["start"]) + "] " + seg["speaker"] + ": " + seg["text"]
            t = self.count(ftext)
            if cur["token_count"] + t > E and cur["segments"]:
                finalize(cur)
                cur = new(seg["start"])
            if t > E:
                for sseg, stext, stok in self.expand(seg):
                    if cur["token_count"] > 0 and cur["token_count"] + stok > E:
                        finalize(cur)
                        cur = new(sseg["start"], sseg, stext, stok)
                    else:
                        append(cur, sseg, stext, stok)
            else:
                append(cur, seg, ftext, t)
        if cur["segments"]:
            finalize(cur)
        for ch in chunks:
            ch["total_chunks"] = len(chunks)
        return chunks
