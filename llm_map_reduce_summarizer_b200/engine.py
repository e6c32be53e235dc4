"""Host side of the path: packing (flat UTF-8 buffer + CSR offsets), the ctypes calls, and dict
materialisation in the reference's output format.  No compute happens here.
"""
import ctypes as C
import math
import os

import numpy as np

from . import _native as N
from . import vocab as _vocab


class TrcError(RuntimeError):
    pass


class PackedBatch:
    """A batch of transcripts in the C ABI's layout (include/trchunk.h: trc_batch)."""

    def __init__(self, text, seg_off, start, end, speaker, doc_off, speakers, originals=None):
        self.text = np.ascontiguousarray(text, dtype=np.uint8)
        self.seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
        self.start = np.ascontiguousarray(start, dtype=np.float64)
        self.end = np.ascontiguousarray(end, dtype=np.float64)
        self.speaker = np.ascontiguousarray(speaker, dtype=np.int32)
        self.doc_off = np.ascontiguousarray(doc_off, dtype=np.int64)
        self.speakers = list(speakers)               # id -> str, sorted, speakers[0] == ""
        self.spk_bytes, self.spk_off = _vocab.pack_strings([s.encode("utf-8") for s in self.speakers])
        self.originals = originals                   # (start objects, end objects) per raw segment, or None

    @property
    def n_seg(self):
        return len(self.start)

    @property
    def n_doc(self):
        return len(self.doc_off) - 1


def _finite_f64(values):
    arr = np.fromiter(map(float, values), dtype=np.float64, count=len(values))
    if arr.size and not np.isfinite(arr).all():
        if np.isnan(arr).any():
            raise ValueError("cannot convert float NaN to integer")
        raise OverflowError("cannot convert float infinity to integer")
    return arr


def pack_documents(docs):
    """list of transcripts (each a list of {start,end,text,speaker} dicts) -> PackedBatch.
    Mirrors the defaults of preprocessor.py:46-50 (`segment.get(key, default)`)."""
    texts, starts, ends, spk_names, doc_off = [], [], [], [], [0]
    for segs in docs:
        texts += [seg.get("text", "") for seg in segs]
        starts += [seg.get("start", 0) for seg in segs]
        ends += [seg.get("end", 0) for seg in segs]
        spk_names += [seg.get("speaker", "") for seg in segs]
        doc_off.append(len(texts))
    for t in texts:
        if type(t) is not str:
            t.strip                                    # same AttributeError as the reference for non-str text
    for s in spk_names:
        if type(s) is not str and not isinstance(s, str):
            raise TypeError("speaker must be a str (the reference's header join fails otherwise)")
    enc = [t.encode("utf-8") for t in texts]          # lone surrogates raise UnicodeEncodeError (unsupported)
    names = sorted(set(spk_names) | {""})
    ids = {s: i for i, s in enumerate(names)}
    buf, off = _vocab.pack_strings(enc)
    st, en = _finite_f64(starts), _finite_f64(ends)
    spk = np.fromiter(map(ids.__getitem__, spk_names), dtype=np.int32, count=len(spk_names))
    return PackedBatch(buf, off, st, en, spk, np.array(doc_off, dtype=np.int64), names, (starts, ends))


class IngestedBatch(PackedBatch):
    """A batch parsed from JSON text by trc_ingest_json (include/trchunk.h).  The arrays are views of the
    library's buffers (freed with this object); the start / end objects json.load would have produced (float,
    int or bool) are only materialised when dicts are built."""

    def __init__(self, lib, handle, view, start_kind, end_kind):
        self._lib, self._handle = lib, handle
        self._originals = None
        n, nd, ns = view.n_seg, view.n_doc, view.n_spk

        def arr(ptr, count, dtype):
            if count == 0 or not ptr:
                return np.zeros(0, dtype=dtype)
            ct = {np.uint8: C.c_uint8, np.int64: C.c_int64, np.float64: C.c_double, np.int32: C.c_int32}[dtype]
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(count,))
        seg_off = arr(view.seg_off, n + 1, np.int64)
        spk_off = arr(view.spk_off, ns + 1, np.int64)
        sb = arr(view.spk_bytes, int(spk_off[-1]), np.uint8).tobytes()
        speakers = [sb[int(spk_off[k]):int(spk_off[k + 1])].decode("utf-8") for k in range(ns)]
        super().__init__(arr(view.text, int(seg_off[-1]), np.uint8), seg_off, arr(view.start, n, np.float64),
                         arr(view.end, n, np.float64), arr(view.speaker, n, np.int32), arr(view.doc_off, nd + 1, np.int64),
                         speakers, None)
        self.start_kind, self.end_kind = arr(start_kind, n, np.uint8), arr(end_kind, n, np.uint8)

    @staticmethod
    def _objects(values, kinds):
        out = values.tolist()
        for k in np.flatnonzero(kinds).tolist():           # ints and bools are the exception in transcripts
            out[k] = int(out[k]) if kinds[k] == 1 else bool(out[k])
        return out

    @property
    def originals(self):
        if self._originals is None:
            self._originals = (self._objects(self.start, self.start_kind), self._objects(self.end, self.end_kind))
        return self._originals

    @originals.setter
    def originals(self, v):
        if v is not None:
            self._originals = v

    def __del__(self):
        try:
            if self._handle:
                self._lib.trc_ingest_free(self._handle)
                self._handle = None
        except Exception:
            pass


def ingest_json(docs, n_threads=None, lib=None):
    """Transcripts as JSON text (list of bytes / str, each {"segments": [...]} as main.py:337-339 loads it) ->
    IngestedBatch, parsed by the library on host threads.  Raises ValueError with the library's message for
    anything json.load or the reference's first use of a field would reject."""
    lib = lib or default_engine().lib
    raw = [d.encode("utf-8") if isinstance(d, str) else bytes(d) for d in docs]
    buf, off = _vocab.pack_strings(raw)
    h = C.c_void_p()
    err = C.create_string_buffer(512)
    nt = n_threads or min(len(raw), os.cpu_count() or 1) or 1
    rc = lib.trc_ingest_json(buf.ctypes.data, off.ctypes.data, len(raw), int(nt), C.byref(h), err, len(err))
    if rc != 0:
        raise ValueError("trc_ingest_json: " + err.value.decode("utf-8", "replace"))
    view = N.Batch()
    sk, ek = C.c_void_p(), C.c_void_p()
    lib.trc_ingest_view(h, C.byref(view), C.byref(sk), C.byref(ek))
    return IngestedBatch(lib, h, view, sk.value, ek.value)


class ProcessedBatch(PackedBatch):
    """Already processed segments (big_chunkeroosky.py:46 accepts any such list): the PIECES are the packed
    segments; `groups` holds the C ABI's extra arrays (include/trchunk.h: trc_batch.already_processed)."""
    groups = None


def pack_processed(segments, for_aggregate=False):
    """One transcript given as processed-segment dicts (the output format of preprocess_transcript,
    preprocessor.py:45-52,203-213) -> ProcessedBatch.  The chunker looks at the components of a segment when
    it has 'segment_timestamps' AND is_combined (big_chunkeroosky.py:281), the interval aggregation whenever
    it has 'segment_timestamps' (preprocessor.py:262): for_aggregate selects the latter."""
    texts, starts, ends, spk_names, group_off, gs, ge, full = [], [], [], [], [0], [], [], []
    for seg in segments:
        sp = seg["speaker"]
        if not isinstance(sp, str):
            raise TypeError("speaker must be a str (the reference's header join fails otherwise)")
        if "segment_timestamps" in seg and (for_aggregate or seg.get("is_combined", False)):
            pieces = seg["segment_timestamps"]
            if len(pieces) < 2:
                raise NotImplementedError("combined segment with fewer than two segment_timestamps entries")
        else:
            pieces = [seg]
        for p in pieces:
            texts.append(p["text"].encode("utf-8"))
            starts.append(p["start"])
            ends.append(p["end"])
            spk_names.append(sp)
        group_off.append(len(texts))
        gs.append(float(seg["start"]))
        ge.append(float(seg["end"]))
        full.append(seg["text"].encode("utf-8"))
    names = sorted(set(spk_names) | {""})
    ids = {s: i for i, s in enumerate(names)}
    buf, off = _vocab.pack_strings(texts)
    st = np.array([float(x) for x in starts], dtype=np.float64)
    en = np.array([float(x) for x in ends], dtype=np.float64)
    for arr in (st, en):
        if arr.size and not np.isfinite(arr).all():
            if np.isnan(arr).any():
                raise ValueError("cannot convert float NaN to integer")
            raise OverflowError("cannot convert float infinity to integer")
    b = ProcessedBatch(buf, off, st, en, np.array([ids[s] for s in spk_names], dtype=np.int32),
                       np.array([0, len(texts)], dtype=np.int64), names, (starts, ends))
    fb, fo = _vocab.pack_strings(full)
    b.groups = {"group_off": np.array(group_off, dtype=np.int64), "group_start": np.array(gs, dtype=np.float64),
                "group_end": np.array(ge, dtype=np.float64), "full_text": fb, "full_off": fo}
    return b


class Engine:
    """One context on one GPU.  Not thread-safe (one engine per thread / per rank)."""

    def __init__(self, device=0, lib_path=None, tokenizer_name="cl100k_base", rank_table=None):
        """rank_table: optional (name, tokens: list[bytes], ranks: list[int]) instead of the table
        `tokenizer_name` resolves to (any byte-pair table whose ranks are its merge priorities)."""
        self.lib = N.load_library(lib_path)
        self._lib_path = lib_path
        h = C.c_void_p()
        rc = self.lib.trc_create(int(device), C.byref(h))
        if rc != 0:
            raise TrcError(
                f"trc_create(device={device}) failed ({N.ERRORS.get(rc, rc)}): a CUDA device of compute capability "
                "10.x (B200, sm_100a) is required; this library has no CPU path")
        self.h = h
        self.device = device
        props = _vocab.load_props()
        self._check(self.lib.trc_load_tables(self.h, props.ctypes.data))
        self.vocab_name, toks, ranks = rank_table if rank_table is not None else _vocab.load_rank_table(tokenizer_name)
        tb, to = _vocab.pack_strings(toks)
        rk = np.array(ranks, dtype=np.int32)
        self.specials = list(_vocab.SPECIALS)
        sb, so = _vocab.pack_strings([s.encode() for s in self.specials])
        self._check(self.lib.trc_load_vocab(self.h, tb.ctypes.data, to.ctypes.data, rk.ctypes.data, len(ranks),
                                            sb.ctypes.data, so.ctypes.data, len(self.specials)))
        self.batch = None
        self.batch_serial = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.trc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            msg = self.lib.trc_last_error(self.h).decode("utf-8", "replace")
            if rc == -4:
                raise NotImplementedError(msg)
            raise TrcError(f"{N.ERRORS.get(rc, rc)}: {msg}")

    # ---- stages ---------------------------------------------------------------------------------------
    def preprocess(self, batch, merge_same_speaker=True, max_segment_duration=120, preserve_timestamps=True,
                   want_merged_text=True, stream=None, device_ptrs=None, seg_len32=None):
        """Stage P on a PackedBatch (host arrays), or on device pointers when `device_ptrs` is given
        (dict with the trc_batch pointer fields as ints).  seg_len32: int32 array of the segments' byte lengths
        to upload instead of the int64 offsets (half the bytes over PCIe)."""
        b = N.Batch()
        if device_ptrs is None:
            b.text, b.seg_off = batch.text.ctypes.data, batch.seg_off.ctypes.data
            if seg_len32 is not None:
                if seg_len32.dtype != np.int32 or seg_len32.shape != (batch.n_seg,):
                    raise ValueError("seg_len32 must be int32[n_seg]")
                b.seg_off, b.seg_len32, b.text_bytes = None, seg_len32.ctypes.data, int(batch.text.size)
            b.start, b.end, b.speaker = batch.start.ctypes.data, batch.end.ctypes.data, batch.speaker.ctypes.data
            b.doc_off = batch.doc_off.ctypes.data
            b.spk_bytes, b.spk_off = batch.spk_bytes.ctypes.data, batch.spk_off.ctypes.data
            b.memory = N.TRC_MEM_HOST
        else:
            for k in ("text", "seg_off", "start", "end", "speaker", "doc_off", "spk_bytes", "spk_off"):
                setattr(b, k, device_ptrs[k])
            b.memory = N.TRC_MEM_DEVICE
        b.n_seg, b.n_doc, b.n_spk = batch.n_seg, batch.n_doc, len(batch.speakers)
        b.merge_same_speaker = 1 if merge_same_speaker else 0
        b.has_max_duration = 0 if max_segment_duration is None else 1
        b.max_duration = float(max_segment_duration) if max_segment_duration is not None else 0.0
        b.preserve_timestamps = 1 if preserve_timestamps else 0
        b.want_merged_text = 1 if want_merged_text else 0
        g = getattr(batch, "groups", None)
        if g is not None:
            if device_ptrs is not None:
                raise ValueError("already-processed batches are passed from host memory")
            b.already_processed = 1
            if g:
                b.group_off, b.n_groups = g["group_off"].ctypes.data, len(g["group_start"])
                b.group_start, b.group_end = g["group_start"].ctypes.data, g["group_end"].ctypes.data
                b.full_text, b.full_off = g["full_text"].ctypes.data, g["full_off"].ctypes.data
        self._keep = (batch, b)
        self._check(self.lib.trc_preprocess(self.h, C.byref(b), stream))
        self.batch = batch
        self.batch_serial += 1
        return self.batch_serial

    def chunk(self, doc_max_tokens, context_tokens=150, add_context=True, stream=None):
        mt = np.ascontiguousarray(doc_max_tokens, dtype=np.int32)
        if mt.shape != (self.batch.n_doc,):
            raise ValueError("doc_max_tokens must have one entry per document")
        self._check(self.lib.trc_chunk(self.h, mt.ctypes.data, int(context_tokens), 1 if add_context else 0, stream))

    def aggregate(self, interval_seconds, stream=None):
        """Fixed time intervals over the resident batch's merged segments (preprocessor.py:217-324)."""
        self._check(self.lib.trc_aggregate(self.h, float(interval_seconds), stream))

    def fetch_aggregated(self):
        return {"agg": self.array("AGG", N.AGG_DTYPE), "mem": self.array("AGG_MEM", N.AGG_MEM_DTYPE),
                "piece": self.array("AGG_PIECE", "<i4"), "text": self.array("AGG_TEXT", "u1"),
                "adoc_off": self.array("AGG_DOC_OFF", "<i8")}

    def assemble_prompts(self, prompt_template, **fields):
        """prompt_template.format(transcript=chunk['text_with_context'], **fields) for every chunk of the resident
        batch (llm_executor.py:187-190), assembled on the device: returns (uint8 buffer, int64 offsets[n_chunks+1]).
        The template is formatted once on the host with a placeholder for the transcript; the device copies the
        literal pieces and the chunk texts."""
        import string
        for _, name, spec, conv in string.Formatter().parse(prompt_template):
            if name is not None and name.split(".")[0].split("[")[0] == "transcript" and (name != "transcript" or spec or conv):
                raise NotImplementedError("only a plain {transcript} field is assembled on the device")
        mark = "\x00\x01trc-transcript\x01\x00"
        if mark in prompt_template or any(mark in str(v) for v in fields.values()):
            raise ValueError("template contains the internal placeholder")
        lits = [x.encode("utf-8") for x in prompt_template.format(transcript=mark, **fields).split(mark)]
        buf, off = _vocab.pack_strings(lits)
        if buf.size == 0:
            buf = np.zeros(1, dtype=np.uint8)
        self._check(self.lib.trc_assemble_prompts(self.h, buf.ctypes.data, off.ctypes.data, len(lits), None))
        return self.array("PROMPTS", "u1"), self.array("PROMPT_OFF", "<i8")

    def counts(self):
        c = np.zeros(N.N_COUNTERS, dtype=np.int64)
        self._check(self.lib.trc_get_counts(self.h, c.ctypes.data))
        return {k: int(c[i]) for i, k in enumerate(N.COUNTERS)}

    def array(self, name, dtype, stream=None):
        nb = C.c_int64()
        self._check(self.lib.trc_get_array(self.h, N.A[name], None, 0, N.TRC_MEM_HOST, C.byref(nb), stream))
        dt = np.dtype(dtype)
        out = np.empty(nb.value // dt.itemsize, dtype=dt)
        if nb.value:
            self._check(self.lib.trc_get_array(self.h, N.A[name], out.ctypes.data, nb.value, N.TRC_MEM_HOST,
                                               C.byref(nb), stream))
        return out

    def array_into(self, name, dst_ptr, dst_bytes, memory=N.TRC_MEM_HOST, stream=None):
        nb = C.c_int64()
        self._check(self.lib.trc_get_array(self.h, N.A[name], dst_ptr, dst_bytes, memory, C.byref(nb), stream))
        return nb.value

    def arrays_into(self, names, dst_ptr, dst_bytes, memory=N.TRC_MEM_HOST, stream=None):
        """Several result arrays with ONE call and one synchronisation (trc_get_arrays).  Returns (offsets, sizes,
        total); with dst_ptr None only the layout is computed."""
        which = np.array([N.A[n] for n in names], dtype=np.int32)
        off = np.zeros(len(names), dtype=np.int64)
        nb = np.zeros(len(names), dtype=np.int64)
        tot = C.c_int64()
        self._check(self.lib.trc_get_arrays(self.h, which.ctypes.data, len(names), dst_ptr, dst_bytes, memory, off.ctypes.data,
                                            nb.ctypes.data, C.byref(tot), stream))
        return off, nb, tot.value

    def format_timestamps(self, seconds):
        """format_timestamp / _format_time of many values with one kernel launch; the resident batch stays resident."""
        x = np.ascontiguousarray(np.array(list(seconds), dtype=np.float64))
        out = np.zeros(24 * max(len(x), 1), dtype=np.uint8)
        self._check(self.lib.trc_format_timestamps(self.h, x.ctypes.data, len(x), out.ctypes.data, N.TRC_MEM_HOST, None))
        raw = out.tobytes()
        return [raw[24 * i + 1:24 * i + 1 + raw[24 * i]].decode("ascii") for i in range(len(x))]

    def stage_ms(self):
        a, b = C.c_float(), C.c_float()
        self._check(self.lib.trc_last_stage_ms(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def launches(self):
        n = C.c_int64()
        self._check(self.lib.trc_kernel_launches(self.h, C.byref(n)))
        return n.value

    def set_profiling(self, on):
        self._check(self.lib.trc_set_profiling(self.h, 1 if on else 0))

    def profile(self):
        """{kernel name: (total ms, launches)} since profiling was switched on."""
        buf = C.create_string_buffer(1 << 16)
        self._check(self.lib.trc_get_profile(self.h, buf, len(buf)))
        out = {}
        for line in buf.value.decode().splitlines():
            name, ms, n = line.split()
            out[name] = (float(ms), int(n))
        return out

    def count_tokens(self, strings, leading_space=False):
        """len(encode(s)) for each str (big_chunkeroosky.py:83) - returns (counts, special_index)."""
        buf, off = _vocab.pack_strings([s.encode("utf-8") for s in strings])
        cnt = np.zeros(len(strings), dtype=np.uint32)
        sp = np.zeros(len(strings), dtype=np.int32)
        self._check(self.lib.trc_count_tokens(self.h, buf.ctypes.data, off.ctypes.data, len(strings),
                                              1 if leading_space else 0, N.TRC_MEM_HOST, cnt.ctypes.data,
                                              sp.ctypes.data, None))
        return cnt, sp

    # ---- materialisation ------------------------------------------------------------------------------
    def fetch_processed(self):
        """Arrays of stage P for the resident batch."""
        r = {"k2r": self.array("K2R", "<i4"), "coff": self.array("COFF", "<i8"), "clean": self.array("CLEAN", "u1"),
             "gstart": self.array("GSTART", "<i4"), "poff": self.array("POFF", "<i8"),
             "merged": self.array("MERGED", "u1"), "gts": self.array("GTS", "u1"), "gtok": self.array("GTOK", "<i8"),
             "gdoc_off": self.array("GDOC_OFF", "<i8"), "ctok": self.array("CTOK", "<u4")}
        return r

    def fetch_chunks(self):
        return {"item_group": self.array("ITEM_GROUP", "<i4"), "item_sub": self.array("ITEM_SUB", "<i4"),
                "item_tok": self.array("ITEM_TOK", "<i4"), "item_len": self.array("ITEM_LEN", "<i8"),
                "item_dst": self.array("ITEM_DST", "<i8"), "subs": self.array("SUBS", N.SUB_DTYPE),
                "chunks": self.array("CHUNKS", N.CHUNK_DTYPE), "spk_list": self.array("SPK_LIST", "<i4"),
                "twc": self.array("TWC", "u1"), "doc_err": self.array("DOC_ERR", "<i4"),
                "cdoc_off": self.array("CDOC_OFF", "<i8"), "sent": self.array("SENT", N.SENT_DTYPE)}


def _text_of(arrays, key):
    """The _Text view of arrays[key], built once per fetched result (documents of a batch share it)."""
    t = arrays.get("_text_" + key)
    if t is None:
        t = arrays["_text_" + key] = _Text(arrays[key])
    return t


class _Text:
    """Slices of a UTF-8 byte buffer as str; ASCII buffers are decoded once."""

    def __init__(self, arr):
        self.b = arr.tobytes()
        self.s = self.b.decode("ascii") if self.b.isascii() else None

    def get(self, a, b):
        return self.s[a:b] if self.s is not None else self.b[a:b].decode("utf-8")


def build_processed(batch, P, doc=0, piece_dicts=None):
    """Processed-segment dicts of document `doc` in the reference's format (preprocessor.py:45-52,203-213).
    piece_dicts: the caller's own processed dicts when the pieces came in already processed; a group of
    one is then returned as that very object (preprocessor.py:181-183)."""
    if batch.originals is None:                               # synthetic batches: the float64 arrays are the times
        batch.originals = (batch.start.tolist(), batch.end.tolist())
    starts, ends = batch.originals
    g0, g1 = int(P["gdoc_off"][doc]), int(P["gdoc_off"][doc + 1])
    gstart = P["gstart"][g0:g1 + 1].tolist()                 # plain lists: element access on numpy arrays is slow
    p0, p1 = (gstart[0], gstart[-1]) if g1 > g0 else (0, 0)
    k2r = P["k2r"][p0:p1].tolist()
    coff = P["coff"][p0:p1 + 1].tolist()
    poff = P["poff"][p0:p1 + 1].tolist()
    clean, merged = _text_of(P, "clean"), _text_of(P, "merged")
    gts_b = P["gts"].tobytes()
    out = []
    names = batch.speakers
    spk = batch.speaker
    cget, mget = clean.get, merged.get
    for g in range(g0, g1):
        a, b = gstart[g - g0] - p0, gstart[g - g0 + 1] - p0
        ra, rb = k2r[a], k2r[b - 1]
        if piece_dicts is not None and b - a == 1:
            out.append(piece_dicts[ra])
            continue
        o = g * 48
        sf = gts_b[o + 1:o + 1 + gts_b[o]].decode("ascii")
        ef = gts_b[o + 25:o + 25 + gts_b[o + 24]].decode("ascii")
        d = {"start": starts[ra], "end": ends[rb], "start_formatted": sf, "end_formatted": ef,
             "speaker": names[spk[ra]], "text": mget(poff[a], poff[b])}
        if b - a > 1:
            d["is_combined"] = True
            d["original_segments"] = b - a
            d["segment_timestamps"] = [{"start": starts[r], "end": ends[r], "text": cget(coff[j], coff[j + 1])}
                                       for j, r in zip(range(a, b), k2r[a:b])]
        out.append(d)
    return out


def build_aggregated(batch, P, AG, processed, interval_seconds, doc=0):
    """Time-interval dicts of document `doc` in the reference's format (preprocessor.py:305-319).  `processed`
    is the document's processed list (its start / end / speaker objects are reused as they are)."""
    g0 = int(P["gdoc_off"][doc])
    a0, a1 = int(AG["adoc_off"][doc]), int(AG["adoc_off"][doc + 1])
    text = _text_of(AG, "text")
    mem = AG["mem"]
    start_time, end_time = processed[0]["start"], processed[-1]["end"]
    out = []
    for rec in AG["agg"][a0:a1]:
        i = int(rec["interval_index"])
        lo = start_time + (i * interval_seconds)                      # :241-242, on the caller's own number types
        hi = min(lo + interval_seconds, end_time)
        if float(lo) != float(rec["start"]) or float(hi) != float(rec["end"]):
            raise NotImplementedError("interval bounds beyond exact fp64 integer arithmetic")
        ts = rec["ts"].tobytes()
        members = mem[int(rec["mem0"]):int(rec["mem1"])]
        segs = [processed[int(m["group"]) - g0] for m in members]
        who = set(processed[g]["speaker"] for g in sorted(int(m["group"]) - g0 for m in members))   # :296 (input order)
        out.append({"start": lo, "end": hi, "start_formatted": ts[1:1 + ts[0]].decode("ascii"),
                    "end_formatted": ts[25:25 + ts[24]].decode("ascii"),
                    "speaker": ", ".join(who) if len(who) > 1 else next(iter(who)),
                    "text": text.get(int(rec["text_off"]), int(rec["text_off"] + rec["text_len"])),
                    "is_aggregated": True, "interval_index": i, "original_segments": len(members),
                    "segment_timestamps": [{"start": s["start"], "end": s["end"], "speaker": s["speaker"],
                                            "text": text.get(int(m["text_off"]), int(m["text_off"] + m["text_len"]))}
                                           for s, m in zip(segs, members)]})
    return out


def special_token_error(engine, k):
    tok = engine.specials[k]
    return ValueError(
        f"Encountered text corresponding to disallowed special token {tok!r}.\n"
        "If you want this text to be encoded as a special token, "
        f"pass it to `allowed_special`, e.g. `allowed_special={{{tok!r}, ...}}`.\n"
        f"If you want this text to be encoded as normal text, disable the check for this token "
        f"by passing `disallowed_special=(enc.special_tokens_set - {{{tok!r}}})`.\n"
        "To disable this check for all special tokens, pass `disallowed_special=()`.\n")


def build_chunks(engine, batch, P, R, processed, doc=0):
    """Chunk dicts of document `doc` (big_chunkeroosky.py:147-195, key order SURVEY.md A.10).
    `processed` is the list returned by build_processed for the same document (objects are aliased)."""
    err = int(R["doc_err"][doc])
    if err > 0:
        raise special_token_error(engine, err - 1)
    if err < 0:
        raise NotImplementedError("document uses a construct outside the supported domain")
    if batch.originals is None:
        batch.originals = (batch.start.tolist(), batch.end.tolist())
    starts, ends = batch.originals
    k2r, gstart = P["k2r"], P["gstart"]
    g0 = int(P["gdoc_off"][doc])
    twc = _text_of(R, "twc")
    ch = R["chunks"]
    subs = R["subs"]
    names = batch.speakers
    c0, c1 = int(R["cdoc_off"][doc]), int(R["cdoc_off"][doc + 1])
    item_group, item_sub, item_dst, item_len = R["item_group"], R["item_sub"], R["item_dst"], R["item_len"]
    sub_cache = {}

    def sub_dict(si, t):
        if si in sub_cache:
            return sub_cache[si]
        s = subs[si]
        g = int(s["group"])
        parent = processed[g - g0]
        text = twc.get(int(item_dst[t]), int(item_dst[t] + item_len[t]))
        st = starts[int(k2r[s["start_src"]])] if s["start_src"] >= 0 else float(s["start"])
        en = ends[int(k2r[s["end_src"]])] if s["end_src"] >= 0 else float(s["end"])
        if s["kind"] == 2:
            d = {"start": st, "end": en, "speaker": "", "text": text, "is_sub_chunk": True, "is_clause": True}
        else:
            d = {"start": st, "end": en, "speaker": parent["speaker"], "text": text, "is_sub_chunk": True,
                 "parent_segment_start": parent["start"], "parent_segment_end": parent["end"]}
        sub_cache[si] = d
        return d

    out = []
    total = c1 - c0
    for c in range(c0, c1):
        r = ch[c]
        if r["err"]:
            raise NotImplementedError("position_percentage outside the exactly formatted range (|x| >= 2**59)")
        segs = []
        for t in range(int(r["item0"]), int(r["item1"])):
            si = int(item_sub[t])
            segs.append(processed[int(item_group[t]) - g0] if si < 0 else sub_dict(si, t))
        o = int(r["twc_off"])
        hdr = int(r["hdr_len"])
        body0 = o + (hdr + 2 if hdr > 0 else 0)          # the header is never empty when add_context is on
        body1 = body0 + int(r["text_len"])
        text = twc.get(body0, body1)
        tw = twc.get(o, body1) if hdr > 0 else text
        start_time = processed[int(r["start_ref"]) - g0]["start"] if r["start_kind"] == 0 else segs[0]["start"]
        end_time = segs[-1]["end"]
        d = {"segments": segs, "text": text, "token_count": int(r["token_count"]), "start_time": start_time,
             "end_time": end_time,
             "speakers": [names[int(i)] for i in R["spk_list"][int(r["item0"]):int(r["item0"]) + int(r["n_speakers"])]],
             "chunk_index": int(r["chunk_index"]), "total_chunks": total,
             "position_percentage": 0 if r["pct_is_int_zero"] else float(r["position_percentage"]),
             "text_with_context": tw}
        out.append(d)
    return out


_ENGINES = {}


def default_engine(tokenizer_name="cl100k_base"):
    """Process-wide engine used by the drop-in modules (device from $TRCHUNK_DEVICE, default 0)."""
    dev = int(os.environ.get("TRCHUNK_DEVICE", "0"))
    key = (dev, tokenizer_name)
    if key not in _ENGINES:
        _ENGINES[key] = Engine(dev, None, tokenizer_name)
    return _ENGINES[key]


def scratch_engine(tokenizer_name="cl100k_base"):
    """A second context on the same device for stand-alone helpers (`clean_text`): what they run does not replace the
    batch that `preprocess_transcript` left resident on the default engine, so a following `chunk_transcript` or
    `aggregate_by_time_interval` does not have to run stage P again."""
    dev = int(os.environ.get("TRCHUNK_DEVICE", "0"))
    key = (dev, tokenizer_name, "scratch")
    main = default_engine(tokenizer_name)
    if key not in _ENGINES or getattr(_ENGINES[key], "_lib_path", None) != getattr(main, "_lib_path", None):
        # (the library of the default engine: the CPU test-suite installs an engine on the g++ test build)
        _ENGINES[key] = Engine(main.device, getattr(main, "_lib_path", None), tokenizer_name)
    return _ENGINES[key]


def set_default_engine(engine, tokenizer_name="cl100k_base"):
    """Install an already constructed engine as the drop-in modules' engine."""
    dev = int(os.environ.get("TRCHUNK_DEVICE", "0"))
    _ENGINES[(dev, tokenizer_name)] = engine
