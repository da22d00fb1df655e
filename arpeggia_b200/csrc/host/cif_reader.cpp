// mmCIF reader for the host side: the `_atom_site` loop of the first data block that has one, as atom records.
// The reference reads PDB or mmCIF by file extension through pdbtbx (src/utils.rs:51-57,
// ReadOptions{only_atomic_coords: true, level: Loose}); pdbtbx's source is not under /root/reference, so the column
// choice below is this repo's statement of it and is UNPINNED: author identifiers where the file
// has them (auth_asym_id, auth_seq_id -- what the PDB-format file of the same entry carries, so both formats give the
// same frame), label identifiers otherwise.
// One pass over a whole-file buffer; tokens are string views into it; numbers via std::from_chars.
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <string_view>
#include <vector>

#include "common.hpp"

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

}  // namespace

extern "C" {

int arph_parse_cif(const char* path, arp_atom_record** out, int64_t* n) {
  *out = nullptr;
  *n = 0;
  FILE* f = std::fopen(path, "rb");
  if (!f) {
    arp::last_error() = std::string("cannot open ") + path;
    return ARP_ERR_IO;
  }
  std::string text;
  {
    char chunk[1 << 16];
    size_t got;
    while ((got = std::fread(chunk, 1, sizeof chunk, f)) > 0) text.append(chunk, got);
  }
  std::fclose(f);

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
    std::vector<std::string_view> row(width);
    while (have) {
      // a row ends the loop when its first token is a keyword or a tag
      if (!T.quoted && (iequals(tok, "loop_") || istarts(tok, "data_") || istarts(tok, "save_") || (!tok.empty() && tok.front() == '_'))) break;
      size_t k = 0;
      for (; k < width && have; ++k) {
        row[k] = tok;
        have = T.next(tok);
      }
      if (k < width) break;  // truncated last row (Loose: dropped)
      auto get = [&](int c) -> std::string_view {
        if (where[c] < 0) return {};
        const std::string_view v = row[where[c]];
        return is_null(v) ? std::string_view{} : v;
      };
      auto first_of = [&](int a, int b) {
        const std::string_view v = get(a);
        return v.empty() ? get(b) : v;
      };
      arp_atom_record r;
      std::memset(&r, 0, sizeof r);
      r.hetero = iequals(get(GROUP), "HETATM") ? 1 : 0;
      number(get(ID), r.serial);
      number(get(MODEL), r.model);
      store(r.name, sizeof r.name, first_of(LABEL_ATOM, AUTH_ATOM));
      store(r.resn, sizeof r.resn, first_of(LABEL_COMP, AUTH_COMP));
      store(r.chain, sizeof r.chain, first_of(AUTH_ASYM, LABEL_ASYM));
      store(r.altloc, sizeof r.altloc, get(ALT));
      store(r.icode, sizeof r.icode, get(INS));
      number(first_of(AUTH_SEQ, LABEL_SEQ), r.resi);
      if (!number(get(X), r.x) || !number(get(Y), r.y) || !number(get(Z), r.z)) {
        arp::last_error() = "mmCIF _atom_site row " + std::to_string(recs.size() + 1) + ": unreadable coordinate";
        return ARP_ERR_IO;
      }
      r.occupancy = 1.0;
      number(get(OCC), r.occupancy);
      std::string sym;
      for (char ch : get(SYMBOL)) sym.push_back((char)std::toupper((unsigned char)ch));
      if (sym.empty()) {  // no type_symbol: leading letters of the atom name, like the fixed-column reader
        for (char ch : std::string_view(r.name, strnlen(r.name, sizeof r.name)).substr(0, 2))
          if (std::isalpha((unsigned char)ch)) sym.push_back((char)std::toupper((unsigned char)ch));
        if (sym.size() == 2 && arp_element_class(sym.c_str()) < 0) sym.resize(1);
      }
      store(r.element, sizeof r.element, sym);
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
