// Ingest: transcripts as JSON text -> the packed batch layout of include/trchunk.h, on the host.
// Replaces json.load(f) + transcript_data.get('segments', []) (/root/reference/main.py:337-339, :133) followed by the
// per-segment segment.get(...) of preprocessor.py:46-50, without building Python objects: one pass over the bytes
// per document (documents in parallel on host threads), strings scanned eight bytes at a time, numbers through
// strtod (correctly rounded, like Python's float()).  Anything json.load would reject, and anything the
// reference's first use of a field would raise on (non-object segment, non-string text/speaker, null or string
// times, NaN/Infinity, lone surrogates) or that the float64 layout cannot hold exactly (integers beyond 2^53), is an error with the byte offset - never a guess.
#pragma once
#include <stdint.h>
#include <errno.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <atomic>
#include <map>
#include <string>
#include <thread>
#include <vector>

struct TrcIngestDoc {
  std::vector<uint8_t> text; std::vector<int64_t> seg_len; std::vector<double> start, end;
  std::vector<uint8_t> start_kind, end_kind;        // 0 float, 1 int, 2 bool (the Python type json.load would give)
  std::vector<std::string> speaker;
  std::string err;
};

struct trc_ingest {
  std::vector<uint8_t> text; std::vector<int64_t> seg_off, doc_off, spk_off; std::vector<double> start, end;
  std::vector<int32_t> speaker; std::vector<uint8_t> start_kind, end_kind, spk_bytes;
  int64_t n_spk = 0;
};

struct TrcJson {
  const uint8_t* s; int64_t n; int64_t i; std::string err;
  bool fail(const char* what) { if (err.empty()) err = std::string(what) + " at byte " + std::to_string(i); return false; }
  void ws() { while (i < n && (s[i] == ' ' || s[i] == '\n' || s[i] == '\r' || s[i] == '\t')) ++i; }
  static void put_utf8(std::vector<uint8_t>* out, uint32_t cp) {
    if (cp < 0x80) out->push_back((uint8_t)cp);
    else if (cp < 0x800) { out->push_back((uint8_t)(0xC0 | (cp >> 6))); out->push_back((uint8_t)(0x80 | (cp & 63))); }
    else if (cp < 0x10000) { out->push_back((uint8_t)(0xE0 | (cp >> 12))); out->push_back((uint8_t)(0x80 | ((cp >> 6) & 63))); out->push_back((uint8_t)(0x80 | (cp & 63))); }
    else { out->push_back((uint8_t)(0xF0 | (cp >> 18))); out->push_back((uint8_t)(0x80 | ((cp >> 12) & 63)));
           out->push_back((uint8_t)(0x80 | ((cp >> 6) & 63))); out->push_back((uint8_t)(0x80 | (cp & 63))); }
  }
  bool hex4(uint32_t* v) {
    if (i + 4 > n) return fail("truncated \\u escape");
    uint32_t x = 0;
    for (int k = 0; k < 4; ++k) {
      uint8_t c = s[i + k]; uint32_t d;
      if (c >= '0' && c <= '9') d = c - '0'; else if (c >= 'a' && c <= 'f') d = c - 'a' + 10; else if (c >= 'A' && c <= 'F') d = c - 'A' + 10;
      else return fail("bad \\u escape");
      x = x * 16 + d;
    }
    i += 4; *v = x; return true;
  }
  // s[i] == '"': the string's content, unescaped to UTF-8, appended to *out (out may be null: validate only)
  bool str(std::vector<uint8_t>* out) {
    ++i;
    for (;;) {
      // plain run: eight bytes at a time until a quote, a backslash, a control byte or a non-ASCII byte shows up
      int64_t j = i;
      while (j + 8 <= n) {
        uint64_t x; memcpy(&x, s + j, 8);
        const uint64_t H = 0x8080808080808080ULL, L = 0x0101010101010101ULL;
        uint64_t q = x ^ (L * '"'), b = x ^ (L * '\\');
        uint64_t hit = ((q - L) & ~q) | ((b - L) & ~b) | ((x - L * 0x20) & ~x) | x;   // zero byte tests (may over-report, never under)
        if (hit & H) break;
        j += 8;
      }
      while (j < n && s[j] != '"' && s[j] != '\\' && s[j] >= 0x20 && s[j] < 0x80) ++j;
      if (out && j > i) out->insert(out->end(), s + i, s + j);
      i = j;
      if (i >= n) return fail("unterminated string");
      uint8_t c = s[i];
      if (c == '"') { ++i; return true; }
      if (c < 0x20) return fail("control character in string");
      if (c >= 0x80) {                                  // UTF-8 sequence: validated (json.load decodes the file as UTF-8 first)
        int len = c >= 0xF0 ? 4 : c >= 0xE0 ? 3 : c >= 0xC2 ? 2 : 0;
        if (!len || c > 0xF4 || i + len > n) return fail("invalid UTF-8");
        uint32_t cp = len == 2 ? (c & 0x1F) : len == 3 ? (c & 0x0F) : (c & 0x07);
        for (int k = 1; k < len; ++k) { uint8_t d = s[i + k]; if ((d & 0xC0) != 0x80) return fail("invalid UTF-8"); cp = (cp << 6) | (d & 63); }
        if ((len == 3 && cp < 0x800) || (len == 4 && (cp < 0x10000 || cp > 0x10FFFF)) || (cp >= 0xD800 && cp < 0xE000)) return fail("invalid UTF-8");
        if (out) out->insert(out->end(), s + i, s + i + len);
        i += len;
        continue;
      }
      ++i;                                              // backslash
      if (i >= n) return fail("unterminated string");
      uint8_t e = s[i++];
      uint32_t cp;
      switch (e) {
        case '"': cp = '"'; break; case '\\': cp = '\\'; break; case '/': cp = '/'; break; case 'b': cp = 8; break;
        case 'f': cp = 12; break; case 'n': cp = 10; break; case 'r': cp = 13; break; case 't': cp = 9; break;
        case 'u': {
          if (!hex4(&cp)) return false;
          if (cp >= 0xD800 && cp < 0xDC00 && i + 6 <= n && s[i] == '\\' && s[i + 1] == 'u') {
            int64_t save = i; uint32_t lo;
            i += 2;
            if (!hex4(&lo)) return false;
            if (lo >= 0xDC00 && lo < 0xE000) cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            else i = save;                              // not a pair: the first one stays a lone surrogate
          }
          if (cp >= 0xD800 && cp < 0xE000) return fail("lone surrogate in string (not encodable as UTF-8)");
          break;
        }
        default: --i; return fail("invalid escape");
      }
      if (out) put_utf8(out, cp);
    }
  }
  // JSON number at s[i] (Python's json: -?(0|[1-9]\d*)(\.\d+)?([eE][+-]?\d+)?); also NaN / Infinity / -Infinity, which
  // json.load accepts and the reference then fails on (int(nan)): an error here.  kind 1 = int (no fraction/exponent).
  bool number(double* v, uint8_t* kind) {
    const int64_t b = i;
    if (i < n && s[i] == '-') ++i;
    if (i < n && (s[i] == 'N' || s[i] == 'I')) return fail("NaN / Infinity as a time value");
    if (i >= n || s[i] < '0' || s[i] > '9') return fail("invalid number");
    if (s[i] == '0') ++i; else while (i < n && s[i] >= '0' && s[i] <= '9') ++i;
    bool is_int = true;
    if (i < n && s[i] == '.') {
      ++i; is_int = false;
      if (i >= n || s[i] < '0' || s[i] > '9') return fail("invalid number");
      while (i < n && s[i] >= '0' && s[i] <= '9') ++i;
    }
    if (i < n && (s[i] == 'e' || s[i] == 'E')) {
      int64_t save = i;
      ++i;
      if (i < n && (s[i] == '+' || s[i] == '-')) ++i;
      if (i < n && s[i] >= '0' && s[i] <= '9') { is_int = false; while (i < n && s[i] >= '0' && s[i] <= '9') ++i; }
      else i = save;                                     // Python's regex leaves a dangling 'e' to the caller (then a syntax error)
    }
    if (!v) return true;
    char buf[64];
    const int64_t len = i - b;
    if (is_int) {
      if (len > 19 + (s[b] == '-')) { i = b; return fail("integer time beyond 2^53 (the packed layout holds float64)"); }
      memcpy(buf, s + b, len); buf[len] = 0;
      errno = 0;
      const long long x = strtoll(buf, nullptr, 10);
      if (errno || x > 9007199254740992LL || x < -9007199254740992LL) { i = b; return fail("integer time beyond 2^53 (the packed layout holds float64)"); }
      *v = (double)x; *kind = 1;                         // float(int): correctly rounded conversion
      return true;
    }
    std::string big;
    const char* p = buf;
    if (len < 63) { memcpy(buf, s + b, len); buf[len] = 0; } else { big.assign((const char*)s + b, len); p = big.c_str(); }
    *v = strtod(p, nullptr);                             // correctly rounded (glibc), as float() of the literal
    if (!(*v - *v == 0.0)) { i = b; return fail("number overflows float64"); }   // Python gives inf, the reference fails on it
    *kind = 0;
    return true;
  }
  bool lit(const char* w) {
    const int64_t L = (int64_t)strlen(w);
    if (i + L > n || memcmp(s + i, w, L) != 0) return fail("invalid literal");
    i += L; return true;
  }
  // any JSON value, validated and skipped (depth limited like CPython's recursion limit would: 1000)
  bool skip(int depth) {
    if (depth > 1000) return fail("nesting too deep");
    ws();
    if (i >= n) return fail("unexpected end");
    const uint8_t c = s[i];
    if (c == '"') return str(nullptr);
    if (c == '{') {
      ++i; ws();
      if (i < n && s[i] == '}') { ++i; return true; }
      for (;;) {
        ws();
        if (i >= n || s[i] != '"') return fail("expected a key");
        if (!str(nullptr)) return false;
        ws();
        if (i >= n || s[i] != ':') return fail("expected ':'");
        ++i;
        if (!skip(depth + 1)) return false;
        ws();
        if (i < n && s[i] == ',') { ++i; continue; }
        if (i < n && s[i] == '}') { ++i; return true; }
        return fail("expected ',' or '}'");
      }
    }
    if (c == '[') {
      ++i; ws();
      if (i < n && s[i] == ']') { ++i; return true; }
      for (;;) {
        if (!skip(depth + 1)) return false;
        ws();
        if (i < n && s[i] == ',') { ++i; continue; }
        if (i < n && s[i] == ']') { ++i; return true; }
        return fail("expected ',' or ']'");
      }
    }
    if (c == 't') return lit("true");
    if (c == 'f') return lit("false");
    if (c == 'n') return lit("null");
    if (c == 'N') return lit("NaN");
    if (c == 'I') return lit("Infinity");
    if (c == '-' && i + 1 < n && s[i + 1] == 'I') { ++i; return lit("Infinity"); }
    return number(nullptr, nullptr);
  }
  // a time value: number, or true/false (Python bool: int(True) == 1, preprocessor.py:101)
  bool time_value(double* v, uint8_t* kind) {
    ws();
    if (i >= n) return fail("unexpected end");
    if (s[i] == 't') { if (!lit("true")) return false; *v = 1.0; *kind = 2; return true; }
    if (s[i] == 'f') { if (!lit("false")) return false; *v = 0.0; *kind = 2; return true; }
    if (s[i] == '-' || (s[i] >= '0' && s[i] <= '9') || s[i] == 'N' || s[i] == 'I') return number(v, kind);
    return fail("start/end must be a number (the reference's int(seconds) fails otherwise)");
  }
  // one element of the segments array (it must be an object: the reference calls segment.get, preprocessor.py:38)
  bool segment(TrcIngestDoc* D) {
    ws();
    if (i >= n || s[i] != '{') return fail("a segment must be an object");
    ++i;
    const size_t t0 = D->text.size();
    double st = 0.0, en = 0.0; uint8_t sk = 1, ek = 1;                 // defaults of segment.get('start', 0) / ('end', 0): the int 0
    std::vector<uint8_t> key, spk, txt;
    bool have_txt = false;
    ws();
    if (i < n && s[i] == '}') ++i;
    else for (;;) {
      ws();
      if (i >= n || s[i] != '"') return fail("expected a key");
      key.clear();
      if (!str(&key)) return false;
      ws();
      if (i >= n || s[i] != ':') return fail("expected ':'");
      ++i;
      const size_t kl = key.size();
      const char* k = (const char*)key.data();
      if (kl == 4 && !memcmp(k, "text", 4)) {                            // a repeated key: the last one wins, like dict
        ws();
        if (i >= n || s[i] != '"') return fail("text must be a string (the reference's .strip() fails otherwise)");
        txt.clear(); have_txt = true;
        if (!str(&txt)) return false;
      } else if (kl == 7 && !memcmp(k, "speaker", 7)) {
        ws();
        if (i >= n || s[i] != '"') return fail("speaker must be a string");
        spk.clear();
        if (!str(&spk)) return false;
      } else if (kl == 5 && !memcmp(k, "start", 5)) { if (!time_value(&st, &sk)) return false; }
      else if (kl == 3 && !memcmp(k, "end", 3)) { if (!time_value(&en, &ek)) return false; }
      else if (!skip(3)) return false;
      ws();
      if (i < n && s[i] == ',') { ++i; continue; }
      if (i < n && s[i] == '}') { ++i; break; }
      return fail("expected ',' or '}'");
    }
    if (have_txt) D->text.insert(D->text.end(), txt.begin(), txt.end());
    D->seg_len.push_back((int64_t)(D->text.size() - t0));
    D->start.push_back(st); D->end.push_back(en); D->start_kind.push_back(sk); D->end_kind.push_back(ek);
    D->speaker.emplace_back((const char*)spk.data(), spk.size());
    return true;
  }
  // the whole document: an object whose 'segments' member (default: none) is an array of segments
  bool document(TrcIngestDoc* D) {
    if (n >= 3 && s[0] == 0xEF && s[1] == 0xBB && s[2] == 0xBF) i = 3;  // UTF-8 signature, as json.load's utf-8-sig
    ws();
    if (i >= n || s[i] != '{') return fail("the transcript must be a JSON object (transcript_data.get, main.py:133)");
    ++i;
    std::vector<uint8_t> key;
    ws();
    if (i < n && s[i] == '}') ++i;
    else for (;;) {
      ws();
      if (i >= n || s[i] != '"') return fail("expected a key");
      key.clear();
      if (!str(&key)) return false;
      ws();
      if (i >= n || s[i] != ':') return fail("expected ':'");
      ++i;
      if (key.size() == 8 && !memcmp(key.data(), "segments", 8)) {
        D->text.clear(); D->seg_len.clear(); D->start.clear(); D->end.clear(); D->start_kind.clear(); D->end_kind.clear(); D->speaker.clear();
        ws();
        if (i + 1 < n && s[i] == '{') {                 // iterating an empty dict yields nothing; a non-empty one yields str keys
          ++i; ws();
          if (i < n && s[i] == '}') { ++i; goto next_member; }
          return fail("'segments' must be an array");
        }
        if (i >= n || s[i] != '[') return fail("'segments' must be an array");
        ++i; ws();
        if (i < n && s[i] == ']') ++i;
        else for (;;) {
          if (!segment(D)) return false;
          ws();
          if (i < n && s[i] == ',') { ++i; continue; }
          if (i < n && s[i] == ']') { ++i; break; }
          return fail("expected ',' or ']'");
        }
      } else if (!skip(1)) return false;
    next_member:
      ws();
      if (i < n && s[i] == ',') { ++i; continue; }
      if (i < n && s[i] == '}') { ++i; break; }
      return fail("expected ',' or '}'");
    }
    ws();
    if (i != n) return fail("extra data after the document");
    return true;
  }
};

// n_doc documents (bytes json[doc_off[x] .. doc_off[x+1])) -> one packed batch.  Returns "" or the first error.
static std::string trc_ingest_run(const uint8_t* json, const int64_t* doc_off, int64_t n_doc, int n_threads, trc_ingest* R) {
  std::vector<TrcIngestDoc> docs((size_t)n_doc);
  std::atomic<int64_t> next(0);
  auto work = [&]() {
    for (;;) {
      const int64_t x = next.fetch_add(1);
      if (x >= n_doc) return;
      TrcJson J; J.s = json + doc_off[x]; J.n = doc_off[x + 1] - doc_off[x]; J.i = 0;
      if (!J.document(&docs[(size_t)x])) docs[(size_t)x].err = "document " + std::to_string(x) + ": " + J.err;
    }
  };
  if (n_threads < 1) n_threads = 1;
  if ((int64_t)n_threads > n_doc) n_threads = (int)(n_doc > 0 ? n_doc : 1);
  std::vector<std::thread> th;
  for (int t = 1; t < n_threads; ++t) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
  for (auto& d : docs) if (!d.err.empty()) return d.err;
  // speaker table: the distinct names and "", in code point (== UTF-8 byte) order - engine.pack_documents' ids
  std::map<std::string, int32_t> ids;
  ids[""] = 0;
  for (auto& d : docs) for (auto& sp : d.speaker) ids.emplace(sp, 0);
  int32_t k = 0;
  R->spk_off.assign(1, 0);
  for (auto& kv : ids) { kv.second = k++; R->spk_bytes.insert(R->spk_bytes.end(), kv.first.begin(), kv.first.end()); R->spk_off.push_back((int64_t)R->spk_bytes.size()); }
  R->n_spk = k;
  size_t nseg = 0, nbytes = 0;
  for (auto& d : docs) { nseg += d.seg_len.size(); nbytes += d.text.size(); }
  R->text.reserve(nbytes); R->seg_off.reserve(nseg + 1); R->start.reserve(nseg); R->end.reserve(nseg); R->speaker.reserve(nseg);
  R->start_kind.reserve(nseg); R->end_kind.reserve(nseg);
  R->seg_off.push_back(0); R->doc_off.assign(1, 0);
  for (auto& d : docs) {
    R->text.insert(R->text.end(), d.text.begin(), d.text.end());
    for (size_t q = 0; q < d.seg_len.size(); ++q) {
      R->seg_off.push_back(R->seg_off.back() + d.seg_len[q]);
      R->speaker.push_back(ids[d.speaker[q]]);
    }
    R->start.insert(R->start.end(), d.start.begin(), d.start.end()); R->end.insert(R->end.end(), d.end.begin(), d.end.end());
    R->start_kind.insert(R->start_kind.end(), d.start_kind.begin(), d.start_kind.end());
    R->end_kind.insert(R->end_kind.end(), d.end_kind.begin(), d.end_kind.end());
    R->doc_off.push_back((int64_t)R->start.size());
    d = TrcIngestDoc();
  }
  return "";
}
