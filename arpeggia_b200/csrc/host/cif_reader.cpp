// mmCIF reader for the host side: the `_atom_site` loop of the first data block that has one, as atom records.
// The reference reads PDB or mmCIF by file extension through pdbtbx (src/utils.rs:51-57,
// ReadOptions{only_atomic_coords: true, level: Loose}); pdbtbx's source is not under /root/reference, so the column
// choice below is this repo's statement of it and is UNPINNED.  As far as pdbtbx 0.12's _atom_site parser is
// remembered it REQUIRES label_asym_id / label_seq_id / label_comp_id / label_atom_id and builds chain ids and residue
// numbers from them, so the default here is the label identifiers (auth_* only where a label value is absent or null,
// e.g. label_seq_id "." of waters).  Entries whose author and label ids differ (6bft: auth chains A,B,C,G,H,L are label
// chains A..F) then give different chain / resi columns than their PDB-format file: that is the reference's behaviour
// too.  arph_parse_cif_ids(path, 1, ...) selects the author identifiers instead (what the PDB-format file carries).
// One pass over a whole-file buffer; tokens are string views into it; numbers via std::from_chars.
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <string_view>
#include <vector>

#include "common.hpp"
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct Tokenizer {
  std::string_view text;
  size_t pos = 0;
  bool line_start = true;  // pos is at the first column of a line
  bool quoted = false;     // the last token was quoted / a text field (so "_x" or "loop_" inside quotes is a value)

  // next token; false at end of input
  bool next(std::string_view& tok) {
    quoted = false;
    while (pos < text.size()) {
      const char ch = text[pos];
      if (ch == '\n') {
        ++pos;
        line_start = true;
      } else if (ch == ' ' || ch == '\t' || ch == '\r') {
        ++pos;
        line_start = false;
      } else if (ch == '#') {  // comment to end of line
        while (pos < text.size() && text[pos] != '\n') ++pos;
      } else {
        break;
      }
    }
    if (pos >= text.size()) return false;
    const char ch = text[pos];
    if (ch == ';' && line_start) {  // text field: up to the next line that starts with ';'
      const size_t begin = pos + 1;
      size_t end = text.find("\n;", begin);
      if (end == std::string_view::npos) end = text.size();
      tok = text.substr(begin, end - begin);
      pos = std::min(text.size(), end + 2);
      line_start = false;
      quoted = true;
      return true;
    }
    line_start = false;
    if (ch == '\'' || ch == '"') {  // quoted: ends at the same quote followed by white space or end of input
      size_t end = pos + 1;
      while (end < text.size()) {
        if (text[end] == ch && (end + 1 == text.size() || text[end + 1] == ' ' || text[end + 1] == '\t' || text[end + 1] == '\n' ||
                                text[end + 1] == '\r'))
          break;
        if (text[end] == '\n') break;  // unterminated: stop at the line end (Loose)
        ++end;
      }
      tok = text.substr(pos + 1, end - pos - 1);
      pos = std::min(text.size(), end + 1);
      quoted = true;
      return true;
    }
    size_t end = pos;
    while (end < text.size() && text[end] != ' ' && text[end] != '\t' && text[end] != '\n' && text[end] != '\r') ++end;
    tok = text.substr(pos, end - pos);
    pos = end;
    return true;
  }
};

bool iequals(std::string_view a, std::string_view b) {
  if (a.size() != b.size()) return false;
  for (size_t i = 0; i < a.size(); ++i)
    if (std::tolower((unsigned char)a[i]) != std::tolower((unsigned char)b[i])) return false;
  return true;
}
bool istarts(std::string_view a, std::string_view prefix) { return a.size() >= prefix.size() && iequals(a.substr(0, prefix.size()), prefix); }
bool is_null(std::string_view v) { return v == "." || v == "?"; }

template <class T>
bool number(std::string_view v, T& out) {
  if (!v.empty() && v.front() == '+') v.remove_prefix(1);
  if (v.empty()) return false;
  auto r = std::from_chars(v.data(), v.data() + v.size(), out);
  return r.ec == std::errc();
}
// Reals: a value written as [-]digits[.digits] with at most 15 digits (every coordinate and occupancy of an archive file)
// is an integer below 2^53 over an exact power of ten, so ONE IEEE division gives the correctly rounded double -- the
// value from_chars returns, at a fifth of its cost.  Anything else (exponents, longer mantissas, inf / nan, trailing
// text) goes to from_chars as before.
bool number(std::string_view v, double& out) {
  static const double kPow10[16] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15};
  if (!v.empty() && v.front() == '+') v.remove_prefix(1);
  if (v.empty()) return false;
  const char* q = v.data();
  const char* e = q + v.size();
  const bool neg = *q == '-';
  if (neg) ++q;
  uint64_t m = 0;
  int digits = 0, frac = 0;
  bool dot = false;
  for (; q < e; ++q) {
    const unsigned d = (unsigned)(*q - '0');
    if (d <= 9u) {
      m = m * 10u + d;
      ++digits;
      frac += dot ? 1 : 0;
    } else if (*q == '.' && !dot) {
      dot = true;
    } else {
      break;
    }
  }
  if (q == e && digits >= 1 && digits <= 15) {
    double r = (double)m;
    if (frac) r /= kPow10[frac];
    out = neg ? -r : r;
    return true;
  }
  auto r = std::from_chars(v.data(), v.data() + v.size(), out);
  return r.ec == std::errc();
}
void store(char* dst, size_t cap, std::string_view v) {
  std::memset(dst, 0, cap);
  std::memcpy(dst, v.data(), std::min(cap, v.size()));
}

enum Col {
  GROUP, ID, SYMBOL, LABEL_ATOM, AUTH_ATOM, ALT, LABEL_COMP, AUTH_COMP, LABEL_ASYM, AUTH_ASYM, LABEL_SEQ, AUTH_SEQ, INS, X, Y, Z, OCC, MODEL, N_COLS
};
const char* const kColName[N_COLS] = {"group_PDB",     "id",           "type_symbol",  "label_atom_id", "auth_atom_id",      "label_alt_id",
                                      "label_comp_id", "auth_comp_id", "label_asym_id", "auth_asym_id",  "label_seq_id",      "auth_seq_id",
                                      "pdbx_PDB_ins_code", "Cartn_x",  "Cartn_y",      "Cartn_z",       "occupancy",         "pdbx_PDB_model_num"};


size_t piece_bytes() {  // text per host thread (ARPH_CIF_PIECE_BYTES: a small value lets the tests cut small files)
  const char* e = std::getenv("ARPH_CIF_PIECE_BYTES");
  const long v = e ? std::atol(e) : 0;
  return v > 0 ? (size_t)v : (size_t)1 << 20;
}

bool is_loop_end(std::string_view tok) {  // an unquoted keyword or tag where a row would start ends the loop
  return iequals(tok, "loop_") || istarts(tok, "data_") || istarts(tok, "save_") || (!tok.empty() && tok.front() == '_');
}

// one _atom_site row -> one record; false when a coordinate does not parse
bool row_to_record(const std::string_view* row, const int* where, bool auth_ids, arp_atom_record& r) {
  auto get = [&](int c) -> std::string_view {
    if (where[c] < 0) return {};
    const std::string_view v = row[where[c]];
    return is_null(v) ? std::string_view{} : v;
  };
  auto first_of = [&](int a, int b) {
    const std::string_view v = get(a);
    return v.empty() ? get(b) : v;
  };
  std::memset(&r, 0, sizeof r);
  r.hetero = iequals(get(GROUP), "HETATM") ? 1 : 0;
  number(get(ID), r.serial);
  number(get(MODEL), r.model);
  store(r.name, sizeof r.name, first_of(LABEL_ATOM, AUTH_ATOM));
  store(r.resn, sizeof r.resn, first_of(LABEL_COMP, AUTH_COMP));
  store(r.chain, sizeof r.chain, auth_ids ? first_of(AUTH_ASYM, LABEL_ASYM) : first_of(LABEL_ASYM, AUTH_ASYM));
  store(r.altloc, sizeof r.altloc, get(ALT));
  store(r.icode, sizeof r.icode, get(INS));
  number(auth_ids ? first_of(AUTH_SEQ, LABEL_SEQ) : first_of(LABEL_SEQ, AUTH_SEQ), r.resi);
  if (!number(get(X), r.x) || !number(get(Y), r.y) || !number(get(Z), r.z)) return false;
  r.occupancy = 1.0;
  number(get(OCC), r.occupancy);
  char sym[4] = {0, 0, 0, 0};
  size_t ns = 0;
  for (char ch : get(SYMBOL))
    if (ns < 3) sym[ns++] = (char)std::toupper((unsigned char)ch);
  if (ns == 0) {  // no type_symbol: leading letters of the atom name, like the fixed-column reader
    for (char ch : std::string_view(r.name, strnlen(r.name, sizeof r.name)).substr(0, 2))
      if (std::isalpha((unsigned char)ch)) sym[ns++] = (char)std::toupper((unsigned char)ch);
    if (ns == 2 && arp_element_class(sym) < 0) sym[1] = 0, ns = 1;
  }
  store(r.element, sizeof r.element, std::string_view(sym, ns));
  return true;
}

// The _atom_site loop the way every archive file writes it -- one row per line -- on all host threads: the text from the
// first value on is cut at line ends, every piece tokenises its lines on its own.  A line that is not exactly one row
// (values wrapped over lines, two rows on a line, a text field) makes the piece give up, and the caller reads the loop
// with the serial tokenizer instead; the first line that starts with a keyword or tag ends the loop.
// The records go straight into the caller's array: a piece of L lines holds at most L rows, so piece t writes from the
// sum of the line counts before it (the few pieces with blank or comment lines are closed up afterwards).
struct CifPiece {
  size_t lines = 0;  // line count of the piece = capacity of its part of the array
  size_t count = 0;  // records written
  bool irregular = false, ended = false;
  int64_t bad_row = -1;  // piece-local index of a row with an unreadable coordinate
};

size_t count_lines(std::string_view text) {
  size_t n = 0;
  const char* p = text.data();
  const char* end = p + text.size();
  while (p < end) {
    const void* q = std::memchr(p, '\n', (size_t)(end - p));
    if (!q) break;
    ++n;
    p = static_cast<const char*>(q) + 1;
  }
  return n + (p < end ? 1 : 0);  // a last line without a line end
}

void parse_piece(std::string_view text, size_t width, const int* where, bool auth_ids, arp_atom_record* dst, CifPiece& P) {
  std::vector<std::string_view> row(width);
  size_t pos = 0;
  while (pos < text.size()) {
    size_t eol = text.find('\n', pos);
    if (eol == std::string_view::npos) eol = text.size();
    const std::string_view line = text.substr(pos, eol - pos);
    pos = eol + 1;
    if (!line.empty() && line.front() == ';') {
      P.irregular = true;
      return;
    }
    size_t k = 0;
    bool first_quoted = false;
    {
      // the plain case -- unquoted values separated by blanks -- in a tight loop; a quote anywhere hands the line to the
      // general tokenizer
      const char* q = line.data();
      const char* e = q + line.size();
      bool quotes = false;
      while (true) {
        while (q < e && (*q == ' ' || *q == '\t' || *q == '\r')) ++q;
        if (q >= e || *q == '#') break;  // end of line / comment to the end of the line
        if (*q == '\'' || *q == '"') {
          quotes = true;
          break;
        }
        const char* b = q;
        while (q < e && *q != ' ' && *q != '\t' && *q != '\r') ++q;
        if (k < width) row[k] = std::string_view(b, (size_t)(q - b));
        ++k;
      }
      if (quotes) {
        Tokenizer T{line};
        T.line_start = false;
        std::string_view tok;
        k = 0;
        while (T.next(tok)) {
          if (k == 0) first_quoted = T.quoted;
          if (k < width) row[k] = tok;
          ++k;
        }
      }
    }
    if (k == 0) continue;  // blank or comment line
    if (!first_quoted && is_loop_end(row[0])) {
      P.ended = true;
      return;
    }
    if (k != width) {
      P.irregular = true;
      return;
    }
    if (!row_to_record(row.data(), where, auth_ids, dst[P.count])) {
      P.bad_row = (int64_t)P.count;
      return;
    }
    P.count += 1;
  }
}

}  // namespace

extern "C" {

int arph_parse_cif_ids(const char* path, int use_auth_ids, arp_atom_record** out, int64_t* n) {
  arp::NvtxRange nvtx_("arph:parse_cif");
  const bool auth_ids = use_auth_ids != 0;
  arp::PhaseClock clk;
  *out = nullptr;
  *n = 0;
  // a regular file is mapped, not copied: its pages come in under the threads that tokenize them; anything else (a pipe)
  // is read in 64 KB pieces
  arp::FileText file;
  if (!file.open(path)) {
    arp::last_error() = std::string("cannot open ") + path;
    return ARP_ERR_IO;
  }
  const std::string_view text = file.view;
  clk.mark("cif read");

  std::vector<arp_atom_record> recs;
  Tokenizer T{text};
  std::string_view tok;
  bool have = T.next(tok);
  bool done = false;
  while (have && !done) {
    if (T.quoted || !iequals(tok, "loop_")) {
      have = T.next(tok);
      continue;
    }
    // a loop: its tags, then rows of values
    std::vector<int> col_of;  // loop column -> Col or -1
    bool atom_site = false;
    while ((have = T.next(tok)) && !T.quoted && !tok.empty() && tok.front() == '_') {
      int c = -1;
      if (istarts(tok, "_atom_site.")) {
        atom_site = true;
        const std::string_view name = tok.substr(11);
        for (int k = 0; k < N_COLS; ++k)
          if (iequals(name, kColName[k])) c = k;
      }
      col_of.push_back(c);
    }
    if (!atom_site || col_of.empty()) continue;  // `tok` is the first value of some other loop: the outer scan skips its values
    int where[N_COLS];
    for (int k = 0; k < N_COLS; ++k) where[k] = -1;
    for (size_t k = 0; k < col_of.size(); ++k)
      if (col_of[k] >= 0) where[col_of[k]] = (int)k;
    if (where[X] < 0 || where[Y] < 0 || where[Z] < 0) {
      arp::last_error() = "mmCIF _atom_site loop without Cartn_x / Cartn_y / Cartn_z";
      return ARP_ERR_IO;
    }
    const size_t width = col_of.size();
    if (have) {  // one row per line? then on all host threads
      const size_t data_begin = (size_t)(tok.data() - text.data()) - (T.quoted ? 1 : 0);
      const std::string_view body = text.substr(data_begin);
      int NT = 1;
#ifdef _OPENMP
      NT = std::max(1, std::min(omp_get_max_threads(), (int)(body.size() / piece_bytes()) + 1));
#endif
      std::vector<size_t> cut((size_t)NT + 1, body.size());
      cut[0] = 0;
      for (int t = 1; t < NT; ++t) {
        size_t p = body.find('\n', std::max(body.size() / (size_t)NT * (size_t)t, cut[(size_t)t - 1]));
        cut[(size_t)t] = p == std::string_view::npos ? body.size() : p + 1;
      }
      std::vector<CifPiece> pieces((size_t)NT);
      const auto piece_text = [&](int t) { return body.substr(cut[(size_t)t], cut[(size_t)t + 1] - cut[(size_t)t]); };
#pragma omp parallel for schedule(static, 1) num_threads(NT)
      for (int t = 0; t < NT; ++t) pieces[(size_t)t].lines = count_lines(piece_text(t));
      std::vector<size_t> first((size_t)NT + 1, 0);
      for (int t = 0; t < NT; ++t) first[(size_t)t + 1] = first[(size_t)t] + pieces[(size_t)t].lines;
      arp_atom_record* arr = (arp_atom_record*)std::malloc(std::max<size_t>(1, first[(size_t)NT]) * sizeof(arp_atom_record));
      if (!arr) {
        arp::last_error() = "out of host memory";
        return ARP_ERR_OOM;
      }
      clk.mark("cif header, cuts, lines");
#pragma omp parallel for schedule(static, 1) num_threads(NT)
      for (int t = 0; t < NT; ++t) parse_piece(piece_text(t), width, where, auth_ids, arr + first[(size_t)t], pieces[(size_t)t]);
      clk.mark("cif parse pieces");
      bool regular = true;
      size_t total = 0;
      for (int t = 0; t < NT; ++t) {
        const CifPiece& pc = pieces[(size_t)t];
        if (pc.irregular) {
          regular = false;
          break;
        }
        if (pc.bad_row >= 0) {
          std::free(arr);
          arp::last_error() = "mmCIF _atom_site row " + std::to_string(total + (size_t)pc.bad_row + 1) + ": unreadable coordinate";
          return ARP_ERR_IO;
        }
        if (pc.count && total != first[(size_t)t]) std::memmove(arr + total, arr + first[(size_t)t], pc.count * sizeof(arp_atom_record));  // close a gap
        total += pc.count;
        if (pc.ended) break;
      }
      if (regular) {  // the first _atom_site loop is the structure
        *n = (int64_t)total;
        arp_atom_record* fit = (arp_atom_record*)std::realloc(arr, std::max<size_t>(1, total) * sizeof(arp_atom_record));
        *out = fit ? fit : arr;
        clk.mark("cif close up");
        return ARP_OK;
      }
      std::free(arr);
    }
    std::vector<std::string_view> row(width);
    while (have) {
      // a row ends the loop when its first token is a keyword or a tag
      if (!T.quoted && is_loop_end(tok)) break;
      size_t k = 0;
      for (; k < width && have; ++k) {
        row[k] = tok;
        have = T.next(tok);
      }
      if (k < width) break;  // truncated last row (Loose: dropped)
      arp_atom_record r;
      if (!row_to_record(row.data(), where, auth_ids, r)) {
        arp::last_error() = "mmCIF _atom_site row " + std::to_string(recs.size() + 1) + ": unreadable coordinate";
        return ARP_ERR_IO;
      }
      recs.push_back(r);
    }
    done = true;  // the first _atom_site loop is the structure
  }
  *n = (int64_t)recs.size();
  *out = (arp_atom_record*)std::malloc(std::max<size_t>(1, recs.size()) * sizeof(arp_atom_record));
  if (!*out) {
    arp::last_error() = "out of host memory";
    return ARP_ERR_OOM;
  }
  if (!recs.empty()) std::memcpy(*out, recs.data(), recs.size() * sizeof(arp_atom_record));
  return ARP_OK;
}

int arph_parse_cif(const char* path, arp_atom_record** out, int64_t* n) { return arph_parse_cif_ids(path, 0, out, n); }

// pdbtbx dispatches on the extension (src/utils.rs:51-57): .cif / .mmcif -> mmCIF, everything else -> PDB
int arph_parse_file(const char* path, arp_atom_record** out, int64_t* n) {
  const std::string p(path ? path : "");
  std::string ext;
  const size_t dot = p.find_last_of('.');
  if (dot != std::string::npos && p.find('/', dot) == std::string::npos)
    for (char ch : p.substr(dot)) ext.push_back((char)std::tolower((unsigned char)ch));
  if (ext == ".cif" || ext == ".mmcif") return arph_parse_cif(path, out, n);
  return arph_parse_pdb(path, out, n);
}
}
