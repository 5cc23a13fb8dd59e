// Host-side tokenisation for the tutoring path: what sits either side of the GPU work.
//   * GPT-2 byte-level BPE encode / decode  - REF/tutoring_server.py:20 (tokenizer.encode) and :30 (tokenizer.decode)
//   * BERT WordPiece encode with [CLS] / [SEP] and truncation to max_length - REF/lms_server.py:98
// The reference calls transformers' GPT2Tokenizer / BertTokenizer, which in transformers 5.x are thin wrappers over the
// Rust `tokenizers` library: BPE + ByteLevel pre-tokenizer / decoder, and WordPiece + BertNormalizer + BertPreTokenizer
// + a "[CLS] $A [SEP]" template.  This file restates those algorithms in C++ behind the C ABI; the Unicode classes and
// the per-code-point normalisation come from unicode_tables.inc, which tools/make_unicode_tables.py obtains by probing
// that library one code point at a time.  tests/test_tokenizer.py checks ids and decoded text against the transformers
// classes on vocabularies trained by oracle/make_vocab.py.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <mutex>
#include <queue>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/tutor_b200.h"
#include "common.cuh"

namespace {

#include "unicode_tables.inc"

// ------------------------------------------------------------------------------------------------ UTF-8, classes
inline bool in_ranges(const uint32_t (*r)[2], int n, uint32_t c) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    const int mid = (lo + hi) >> 1;
    if (c < r[mid][0]) hi = mid - 1;
    else if (c > r[mid][1]) lo = mid + 1;
    else return true;
  }
  return false;
}

// next code point of valid UTF-8 (callers encode from str, so input is valid; a stray byte is passed through as itself)
inline uint32_t next_cp(const unsigned char* s, size_t n, size_t& i) {
  const unsigned char b = s[i];
  if (b < 0x80) { ++i; return b; }
  int len = (b >= 0xF0) ? 4 : (b >= 0xE0) ? 3 : (b >= 0xC0) ? 2 : 1;
  if (len == 1 || i + len > n) { ++i; return b; }
  uint32_t c = b & (0xFF >> (len + 1));
  for (int k = 1; k < len; ++k) c = (c << 6) | (s[i + k] & 0x3F);
  i += len;
  return c;
}
inline void put_cp(std::string& out, uint32_t c) {
  if (c < 0x80) out.push_back((char)c);
  else if (c < 0x800) { out.push_back((char)(0xC0 | (c >> 6))); out.push_back((char)(0x80 | (c & 0x3F))); }
  else if (c < 0x10000) {
    out.push_back((char)(0xE0 | (c >> 12))); out.push_back((char)(0x80 | ((c >> 6) & 0x3F))); out.push_back((char)(0x80 | (c & 0x3F)));
  } else {
    out.push_back((char)(0xF0 | (c >> 18))); out.push_back((char)(0x80 | ((c >> 12) & 0x3F)));
    out.push_back((char)(0x80 | ((c >> 6) & 0x3F))); out.push_back((char)(0x80 | (c & 0x3F)));
  }
}
std::vector<uint32_t> to_cps(const char* s, size_t n) {
  std::vector<uint32_t> v;
  v.reserve(n);
  for (size_t i = 0; i < n;) v.push_back(next_cp(reinterpret_cast<const unsigned char*>(s), n, i));
  return v;
}

int read_file(const char* path, std::string& out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) { tutor_set_error("tokenizer: cannot open %s", path); return TUTOR_ERR_ARG; }
  std::stringstream ss;
  ss << f.rdbuf();
  out = ss.str();
  return TUTOR_OK;
}

// positions of the literal special tokens in a text: leftmost match, longest among those starting at the same place
struct Cut { size_t begin, end; int id; };
std::vector<Cut> find_specials(const std::string& text, const std::vector<std::pair<std::string, int>>& specials) {
  std::vector<Cut> cuts;
  size_t i = 0;
  while (i < text.size()) {
    size_t best = std::string::npos, best_len = 0;
    int best_id = -1;
    for (const auto& sp : specials) {
      const size_t p = text.find(sp.first, i);
      if (p == std::string::npos) continue;
      if (p < best || (p == best && sp.first.size() > best_len)) { best = p; best_len = sp.first.size(); best_id = sp.second; }
    }
    if (best == std::string::npos) break;
    cuts.push_back(Cut{best, best + best_len, best_id});
    i = best + best_len;
  }
  return cuts;
}

// ------------------------------------------------------------------------------------------------ GPT-2 BPE
struct PairHash {
  size_t operator()(const std::pair<int, int>& p) const { return (size_t)p.first * 1000003u ^ (size_t)p.second; }
};

}  // namespace

struct tutor_bpe {
  std::unordered_map<std::string, int> vocab;
  std::vector<std::string> id_to_token;
  std::unordered_map<std::pair<int, int>, std::pair<int, int>, PairHash> merges;  // (a, b) -> (rank, merged id)
  std::string byte_to_uni[256];           // byte -> its printable stand-in, UTF-8
  std::unordered_map<uint32_t, int> uni_to_byte;
  int byte_id[256];                       // vocabulary id of every single-byte token
  std::vector<std::pair<std::string, int>> specials;
  int eos_id = -1;
  std::unordered_map<std::string, std::vector<int>> cache;  // pre-token -> ids
  std::mutex mu;
};

namespace {

// JSON object of string -> integer (vocab.json): a scanner for exactly that shape
int parse_vocab_json(const std::string& js, std::unordered_map<std::string, int>& vocab) {
  size_t i = 0;
  const size_t n = js.size();
  auto skip_ws = [&] { while (i < n && (js[i] == ' ' || js[i] == '\n' || js[i] == '\r' || js[i] == '\t')) ++i; };
  auto hex4 = [&](uint32_t& v) {
    if (i + 4 > n) return false;
    v = 0;
    for (int k = 0; k < 4; ++k) {
      const char c = js[i++];
      v <<= 4;
      if (c >= '0' && c <= '9') v |= c - '0';
      else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
      else return false;
    }
    return true;
  };
  skip_ws();
  if (i >= n || js[i] != '{') return -1;
  ++i;
  while (true) {
    skip_ws();
    if (i < n && js[i] == '}') return 0;
    if (i >= n || js[i] != '"') return -1;
    ++i;
    std::string key;
    while (i < n && js[i] != '"') {
      if (js[i] == '\\') {
        if (++i >= n) return -1;
        const char e = js[i++];
        switch (e) {
          case 'n': key.push_back('\n'); break;
          case 't': key.push_back('\t'); break;
          case 'r': key.push_back('\r'); break;
          case 'b': key.push_back('\b'); break;
          case 'f': key.push_back('\f'); break;
          case 'u': {
            uint32_t u;
            if (!hex4(u)) return -1;
            if (u >= 0xD800 && u <= 0xDBFF && i + 6 <= n && js[i] == '\\' && js[i + 1] == 'u') {
              i += 2;
              uint32_t lo;
              if (!hex4(lo)) return -1;
              u = 0x10000 + ((u - 0xD800) << 10) + (lo - 0xDC00);
            }
            put_cp(key, u);
            break;
          }
          default: key.push_back(e);  // \" \\ \/
        }
      } else {
        key.push_back(js[i++]);
      }
    }
    if (i >= n) return -1;
    ++i;
    skip_ws();
    if (i >= n || js[i] != ':') return -1;
    ++i;
    skip_ws();
    bool neg = false;
    if (i < n && js[i] == '-') { neg = true; ++i; }
    if (i >= n || js[i] < '0' || js[i] > '9') return -1;
    long v = 0;
    while (i < n && js[i] >= '0' && js[i] <= '9') v = v * 10 + (js[i++] - '0');
    vocab[key] = (int)(neg ? -v : v);
    skip_ws();
    if (i < n && js[i] == ',') { ++i; continue; }
    if (i < n && js[i] == '}') return 0;
    return -1;
  }
}

enum { G_OTHER = 0, G_LETTER = 1, G_NUMBER = 2, G_SPACE = 3 };
inline int gpt2_class(uint32_t c) {
  if (c < 0x80) {  // the common case without a search
    if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z')) return G_LETTER;
    if (c >= '0' && c <= '9') return G_NUMBER;
    if (c == ' ' || (c >= 9 && c <= 13)) return G_SPACE;
    if (c >= 0x1C && c <= 0x1F) return in_ranges(kGpt2Space, kGpt2Space_n, c) ? G_SPACE : G_OTHER;
    return G_OTHER;
  }
  if (in_ranges(kGpt2Letter, kGpt2Letter_n, c)) return G_LETTER;
  if (in_ranges(kGpt2Number, kGpt2Number_n, c)) return G_NUMBER;
  if (in_ranges(kGpt2Space, kGpt2Space_n, c)) return G_SPACE;
  return G_OTHER;
}

// The GPT-2 pattern  's|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+  over code points:
// returns the end of the token that starts at i (alternatives in order, each greedy, as a backtracking engine would).
size_t gpt2_next_token(const std::vector<uint32_t>& t, size_t i) {
  const size_t n = t.size();
  if (t[i] == '\'' && i + 1 < n) {
    const uint32_t a = t[i + 1], b = i + 2 < n ? t[i + 2] : 0;
    if (a == 's' || a == 't' || a == 'm' || a == 'd') return i + 2;
    if ((a == 'r' && b == 'e') || (a == 'v' && b == 'e') || (a == 'l' && b == 'l')) return i + 3;
  }
  const size_t j = (t[i] == ' ' && i + 1 < n) ? i + 1 : i;  // where the run starts if the optional space is taken
  const int cls = gpt2_class(t[j]);
  if (cls != G_SPACE) {  // ` ?X+` with X the class of t[j] (letters, numbers, or anything else that is not a space)
    size_t k = j;
    while (k < n && gpt2_class(t[k]) == cls) ++k;
    return k;
  }
  // t[i] is white space and what follows the optional space is not the start of a run: \s+(?!\S), else \s+
  size_t k = i;
  while (k < n && gpt2_class(t[k]) == G_SPACE) ++k;
  if (k == n || k - i == 1) return k;  // the whole run at the end of the text, or a single white-space character
  return k - 1;                        // leave the last one to lead the next token
}

// one pre-token (already in the byte-level alphabet, as vocabulary ids of its characters) -> merged ids.
// Same procedure as the library's Word::merge_all: a heap of candidate merges ordered by (rank, position); a popped
// entry is applied only if both symbols are still alive and still form the pair it was created for.
void bpe_merge(const tutor_bpe& tk, std::vector<int>& ids) {
  struct Sym { int id, prev, next; bool alive; };
  struct Cand { int rank, pos, new_id; };
  struct Cmp { bool operator()(const Cand& a, const Cand& b) const { return a.rank != b.rank ? a.rank > b.rank : a.pos > b.pos; } };
  const int n = (int)ids.size();
  if (n < 2) return;
  std::vector<Sym> s(n);
  for (int i = 0; i < n; ++i) s[i] = Sym{ids[i], i - 1, i + 1 < n ? i + 1 : -1, true};
  std::priority_queue<Cand, std::vector<Cand>, Cmp> heap;
  auto push = [&](int pos) {
    const int nx = s[pos].next;
    if (nx < 0) return;
    auto it = tk.merges.find({s[pos].id, s[nx].id});
    if (it != tk.merges.end()) heap.push(Cand{it->second.first, pos, it->second.second});
  };
  for (int i = 0; i + 1 < n; ++i) push(i);
  while (!heap.empty()) {
    const Cand c = heap.top();
    heap.pop();
    if (!s[c.pos].alive) continue;
    const int nx = s[c.pos].next;
    if (nx < 0) continue;
    auto it = tk.merges.find({s[c.pos].id, s[nx].id});
    if (it == tk.merges.end() || it->second.second != c.new_id) continue;
    s[c.pos].id = c.new_id;
    s[c.pos].next = s[nx].next;
    s[nx].alive = false;
    if (s[nx].next >= 0) s[s[nx].next].prev = c.pos;
    if (s[c.pos].prev >= 0) push(s[c.pos].prev);
    push(c.pos);
  }
  ids.clear();
  for (int i = 0; i >= 0; i = s[i].next) ids.push_back(s[i].id);
}

void bpe_encode_segment(tutor_bpe* tk, const char* text, size_t nbytes, std::vector<int>& out) {
  const std::vector<uint32_t> cps = to_cps(text, nbytes);
  std::string piece;
  std::vector<int> ids;
  for (size_t i = 0; i < cps.size();) {
    const size_t e = gpt2_next_token(cps, i);
    piece.clear();
    for (size_t k = i; k < e; ++k) put_cp(piece, cps[k]);
    i = e;
    auto it = tk->cache.find(piece);
    if (it != tk->cache.end()) {
      out.insert(out.end(), it->second.begin(), it->second.end());
      continue;
    }
    ids.clear();
    for (unsigned char b : piece) ids.push_back(tk->byte_id[b]);
    bpe_merge(*tk, ids);
    if (tk->cache.size() < 200000) tk->cache.emplace(piece, ids);
    out.insert(out.end(), ids.begin(), ids.end());
  }
}

}  // namespace

extern "C" {

int tutor_bpe_create(const char* vocab_json_path, const char* merges_txt_path, tutor_bpe_t** out) {
  ARG_CHECK(vocab_json_path && merges_txt_path && out, "bpe_create: null argument");
  std::string js, mg;
  TRY(read_file(vocab_json_path, js));
  TRY(read_file(merges_txt_path, mg));
  tutor_bpe* tk = new tutor_bpe();
  if (parse_vocab_json(js, tk->vocab) != 0 || tk->vocab.empty()) {
    delete tk;
    tutor_set_error("bpe_create: %s is not a JSON object of token -> id", vocab_json_path);
    return TUTOR_ERR_ARG;
  }
  // bytes_to_unicode: printable Latin-1 bytes stand for themselves, the other 68 are moved to U+0100..
  int moved = 0;
  for (int b = 0; b < 256; ++b) {
    const bool keep = (b >= 33 && b <= 126) || (b >= 161 && b <= 172) || (b >= 174 && b <= 255);
    const uint32_t u = keep ? (uint32_t)b : 256u + (uint32_t)moved++;
    put_cp(tk->byte_to_uni[b], u);
    tk->uni_to_byte[u] = b;
  }
  int max_id = -1;
  for (const auto& kv : tk->vocab) max_id = std::max(max_id, kv.second);
  // <|endoftext|> is unk = bos = eos (GPT2Tokenizer's defaults); a vocabulary without it gets it appended, like the library does
  const std::string eot = "<|endoftext|>";
  auto it = tk->vocab.find(eot);
  if (it == tk->vocab.end()) { tk->vocab[eot] = ++max_id; it = tk->vocab.find(eot); }
  tk->eos_id = it->second;
  tk->specials.emplace_back(eot, tk->eos_id);
  tk->id_to_token.assign((size_t)max_id + 1, std::string());
  for (const auto& kv : tk->vocab)
    if (kv.second >= 0) tk->id_to_token[kv.second] = kv.first;
  for (int b = 0; b < 256; ++b) {
    auto bi = tk->vocab.find(tk->byte_to_uni[b]);
    if (bi == tk->vocab.end()) {
      delete tk;
      tutor_set_error("bpe_create: the vocabulary lacks the single-byte token for byte %d (not a byte-level BPE vocabulary)", b);
      return TUTOR_ERR_ARG;
    }
    tk->byte_id[b] = bi->second;
  }
  int rank = 0;
  size_t pos = 0;
  while (pos < mg.size()) {
    size_t eol = mg.find('\n', pos);
    if (eol == std::string::npos) eol = mg.size();
    std::string line = mg.substr(pos, eol - pos);
    pos = eol + 1;
    if (!line.empty() && line.back() == '\r') line.pop_back();
    if (line.empty() || line.rfind("#version", 0) == 0) continue;
    const size_t sp = line.find(' ');
    if (sp == std::string::npos) continue;
    const std::string a = line.substr(0, sp), b = line.substr(sp + 1);
    auto ia = tk->vocab.find(a), ib = tk->vocab.find(b), ic = tk->vocab.find(a + b);
    if (ia != tk->vocab.end() && ib != tk->vocab.end() && ic != tk->vocab.end())
      tk->merges.emplace(std::make_pair(ia->second, ib->second), std::make_pair(rank, ic->second));
    ++rank;
  }
  *out = tk;
  return TUTOR_OK;
}

void tutor_bpe_destroy(tutor_bpe_t* tk) { delete tk; }

int tutor_bpe_eos_id(tutor_bpe_t* tk) { return tk ? tk->eos_id : -1; }
int tutor_bpe_vocab_size(tutor_bpe_t* tk) { return tk ? (int)tk->id_to_token.size() : 0; }

int tutor_bpe_encode(tutor_bpe_t* tk, const char* utf8, int64_t nbytes, int32_t* ids_out, int capacity, int* n_out) {
  ARG_CHECK(tk && (utf8 || nbytes == 0) && n_out && nbytes >= 0, "bpe_encode: bad argument");
  std::lock_guard<std::mutex> lock(tk->mu);
  std::vector<int> ids;
  const std::string text(utf8 ? utf8 : "", (size_t)nbytes);
  size_t at = 0;
  for (const Cut& c : find_specials(text, tk->specials)) {  // the literal special token is one id, never split
    if (c.begin > at) bpe_encode_segment(tk, text.data() + at, c.begin - at, ids);
    ids.push_back(c.id);
    at = c.end;
  }
  if (at < text.size()) bpe_encode_segment(tk, text.data() + at, text.size() - at, ids);
  *n_out = (int)ids.size();
  ARG_CHECK(ids_out || ids.empty(), "bpe_encode: null output");
  ARG_CHECK((int)ids.size() <= capacity, "bpe_encode: %d ids do not fit the output capacity %d", (int)ids.size(), capacity);
  for (size_t i = 0; i < ids.size(); ++i) ids_out[i] = ids[i];
  return TUTOR_OK;
}

int tutor_bpe_decode(tutor_bpe_t* tk, const int32_t* ids, int n, int skip_special_tokens, char* out, int capacity,
                     int* nbytes_out) {
  ARG_CHECK(tk && (ids || n == 0) && nbytes_out && n >= 0, "bpe_decode: bad argument");
  std::string bytes;
  for (int i = 0; i < n; ++i) {
    const int id = ids[i];
    ARG_CHECK(id >= 0 && id < (int)tk->id_to_token.size(), "bpe_decode: id %d outside the vocabulary", id);
    if (id == tk->eos_id) {  // an added token: not part of the byte-level alphabet
      if (!skip_special_tokens) bytes += tk->specials[0].first;
      continue;
    }
    const std::string& tok = tk->id_to_token[id];
    for (size_t k = 0; k < tok.size();) {
      const uint32_t c = next_cp(reinterpret_cast<const unsigned char*>(tok.data()), tok.size(), k);
      auto it = tk->uni_to_byte.find(c);
      if (it != tk->uni_to_byte.end()) bytes.push_back((char)it->second);
      else put_cp(bytes, c);
    }
  }
  *nbytes_out = (int)bytes.size();
  ARG_CHECK((int)bytes.size() <= capacity, "bpe_decode: %d bytes do not fit the output capacity %d", (int)bytes.size(), capacity);
  if (!bytes.empty()) memcpy(out, bytes.data(), bytes.size());
  return TUTOR_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------ BERT WordPiece
struct tutor_wordpiece {
  std::unordered_map<std::string, int> vocab;
  int unk = -1, cls = -1, sep = -1, pad = -1, mask = -1;
  std::vector<std::pair<std::string, int>> specials;
  size_t max_token_bytes = 0;
};

namespace {

// BertNormalizer on one code point -> appended to out (code points)
inline void bert_normalize_cp(uint32_t c, std::vector<uint32_t>& out) {
  if (c < 0x80) {  // ASCII without a search: controls dropped, white space -> ' ', upper -> lower
    if (c == '\t' || c == '\n' || c == '\r' || c == ' ') { out.push_back(' '); return; }
    if (c < 0x20 || c == 0x7F) return;
    out.push_back((c >= 'A' && c <= 'Z') ? c + 32 : c);
    return;
  }
  if (in_ranges(kBertRemove, kBertRemove_n, c)) return;
  if (in_ranges(kBertToSpace, kBertToSpace_n, c)) { out.push_back(' '); return; }
  if (in_ranges(kBertCjk, kBertCjk_n, c)) { out.push_back(' '); out.push_back(c); out.push_back(' '); return; }
  if (c >= 0xAC00 && c <= 0xD7A3) {  // Hangul syllable -> its jamo (canonical decomposition)
    const uint32_t s = c - 0xAC00, t = 0x11A7 + s % 28;
    out.push_back(0x1100 + s / 588);
    out.push_back(0x1161 + (s % 588) / 28);
    if (t != 0x11A7) out.push_back(t);
    return;
  }
  const uint32_t* k = std::lower_bound(kBertReplKey, kBertReplKey + kBertRepl_n, c);
  if (k != kBertReplKey + kBertRepl_n && *k == c) {
    const size_t idx = (size_t)(k - kBertReplKey);
    for (uint32_t o = kBertReplOff[idx]; o < kBertReplOff[idx + 1]; ++o) out.push_back(kBertReplData[o]);
    return;
  }
  out.push_back(c);
}
inline bool bert_is_space(uint32_t c) { return c == ' ' || (c >= 0x80 && in_ranges(kBertWhitespace, kBertWhitespace_n, c)) || (c < 0x80 && c >= 9 && c <= 13); }
inline bool bert_is_punct(uint32_t c) {
  if (c < 0x80) return (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126);
  return in_ranges(kBertPunct, kBertPunct_n, c);
}

// greedy longest-match-first word pieces of one word (code points)
void wordpiece_word(const tutor_wordpiece& tk, const uint32_t* w, size_t n, std::vector<int>& out) {
  if (n > 100) { out.push_back(tk.unk); return; }  // max_input_chars_per_word
  std::vector<size_t> off(n + 1, 0);
  std::string utf8;
  for (size_t i = 0; i < n; ++i) { put_cp(utf8, w[i]); off[i + 1] = utf8.size(); }
  const size_t first = out.size();
  std::string cand;
  size_t start = 0;
  while (start < n) {
    size_t end = n;
    int id = -1;
    while (end > start) {
      cand.assign(start ? "##" : "");
      cand.append(utf8, off[start], off[end] - off[start]);
      if (cand.size() <= tk.max_token_bytes) {
        auto it = tk.vocab.find(cand);
        if (it != tk.vocab.end()) { id = it->second; break; }
      }
      --end;
    }
    if (id < 0) {  // a stretch with no piece in the vocabulary: the whole word is unknown
      out.resize(first);
      out.push_back(tk.unk);
      return;
    }
    out.push_back(id);
    start = end;
  }
}

void wordpiece_segment(const tutor_wordpiece& tk, const char* text, size_t nbytes, std::vector<int>& out) {
  std::vector<uint32_t> norm;
  norm.reserve(nbytes + 8);
  const unsigned char* s = reinterpret_cast<const unsigned char*>(text);
  for (size_t i = 0; i < nbytes;) bert_normalize_cp(next_cp(s, nbytes, i), norm);
  size_t i = 0;
  const size_t n = norm.size();
  while (i < n) {
    if (bert_is_space(norm[i])) { ++i; continue; }
    if (bert_is_punct(norm[i])) { wordpiece_word(tk, &norm[i], 1, out); ++i; continue; }
    size_t e = i;
    while (e < n && !bert_is_space(norm[e]) && !bert_is_punct(norm[e])) ++e;
    wordpiece_word(tk, &norm[i], e - i, out);
    i = e;
  }
}

}  // namespace

extern "C" {

int tutor_wordpiece_create(const char* vocab_txt_path, tutor_wordpiece_t** out) {
  ARG_CHECK(vocab_txt_path && out, "wordpiece_create: null argument");
  std::string txt;
  TRY(read_file(vocab_txt_path, txt));
  tutor_wordpiece* tk = new tutor_wordpiece();
  int id = 0;
  size_t pos = 0;
  while (pos < txt.size()) {
    size_t eol = txt.find('\n', pos);
    if (eol == std::string::npos) eol = txt.size();
    std::string line = txt.substr(pos, eol - pos);
    pos = eol + 1;
    if (!line.empty() && line.back() == '\r') line.pop_back();
    tk->vocab.emplace(line, id++);  // line number = id (an empty line still takes one)
    tk->max_token_bytes = std::max(tk->max_token_bytes, line.size());
  }
  struct { const char* s; int* dst; } sp[] = {{"[UNK]", &tk->unk}, {"[CLS]", &tk->cls}, {"[SEP]", &tk->sep}, {"[PAD]", &tk->pad},
                                              {"[MASK]", &tk->mask}};
  for (auto& e : sp) {
    auto it = tk->vocab.find(e.s);
    if (it == tk->vocab.end()) {
      delete tk;
      tutor_set_error("wordpiece_create: %s lacks the special token %s", vocab_txt_path, e.s);
      return TUTOR_ERR_ARG;
    }
    *e.dst = it->second;
    tk->specials.emplace_back(e.s, it->second);
  }
  *out = tk;
  return TUTOR_OK;
}

void tutor_wordpiece_destroy(tutor_wordpiece_t* tk) { delete tk; }

int tutor_wordpiece_encode(tutor_wordpiece_t* tk, const char* utf8, int64_t nbytes, int max_length, int32_t* ids_out,
                           int capacity, int* n_out) {
  ARG_CHECK(tk && (utf8 || nbytes == 0) && n_out && nbytes >= 0, "wordpiece_encode: bad argument");
  ARG_CHECK(max_length <= 0 || max_length >= 2, "wordpiece_encode: max_length %d leaves no room for [CLS] and [SEP]", max_length);
  std::vector<int> ids;
  const std::string text(utf8 ? utf8 : "", (size_t)nbytes);
  size_t at = 0;
  for (const Cut& c : find_specials(text, tk->specials)) {  // literal [SEP] etc. in the text are the special ids
    if (c.begin > at) wordpiece_segment(*tk, text.data() + at, c.begin - at, ids);
    ids.push_back(c.id);
    at = c.end;
  }
  if (at < text.size()) wordpiece_segment(*tk, text.data() + at, text.size() - at, ids);
  if (max_length > 0 && (int)ids.size() > max_length - 2) ids.resize((size_t)max_length - 2);  // truncation=True
  *n_out = (int)ids.size() + 2;
  ARG_CHECK(ids_out && *n_out <= capacity, "wordpiece_encode: %d ids do not fit the output capacity %d", *n_out, capacity);
  ids_out[0] = tk->cls;
  for (size_t i = 0; i < ids.size(); ++i) ids_out[i + 1] = ids[i];
  ids_out[ids.size() + 1] = tk->sep;
  return TUTOR_OK;
}

}  // extern "C"
