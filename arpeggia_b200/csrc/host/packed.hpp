// Packed structure: device SoA + decode dictionaries (the thing behind the opaque arph_packed).
#pragma once
#include <set>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.hpp"

namespace arp {

struct Packed {
  // device SoA (arp_structure points into these)
  RawVec<double> x, y, z;  // (the per-atom arrays: resized once, filled by every host thread -- RawVec, common.hpp)
  std::vector<double> hx, hy, hz, ring_c, ring_n;
  RawVec<uint32_t> attr, chain_key, res_ordinal, res_id;
  std::vector<uint32_t> res_h_begin, ring_attr, ring_chain_key, ring_res_ordinal;
  std::vector<uint32_t> res_chain_key, res_res_ordinal;  // per residue (compact form of chain_key / res_ordinal)
  std::vector<uint32_t> res_atom_begin;                  // CSR of the residues' atoms (compact form of res_id)
  std::vector<int32_t> res_cb, res_sg;
  RawVec<uint64_t> xyz21;                                // 21-bit lattice words (arp_structure.xyz_milli21; empty when they do not apply)
  RawVec<uint16_t> attr16;                               // attr as 16-bit words
  RawVec<int32_t> x_milli, y_milli, z_milli;        // on-lattice form of x / y / z (empty when a coordinate is off the lattice)
  std::vector<uint32_t> res_tag;                         // per residue: |tag difference| <= 1 <=> same chain, ordinals <= 1 apart
  std::vector<uint32_t> atom_gid, ring_gid;              // ids in a larger structure (arph_packed_set_global_ids); empty = none
  std::vector<std::string> chain_by_rank;                // chain ids in byte order = the low 24 bits of chain_key
  arp_structure s{};

  // decode side: one ResConf per (residue, conformer) = one ResidueId of the reference (residues.rs:7-21)
  struct ResConf {
    int32_t chain, resn, icode, altloc;  // string ids
    int64_t resi, model_serial;
    bool has_sc;
    Plane sc;  // build_sc_plane_positions, complex.rs:486-514
  };
  RawVec<ResConf> resconfs;
  RawVec<uint32_t> atom_rc;     // per atom -> resconfs
  RawVec<int32_t> atom_name;    // per atom -> strings
  RawVec<int64_t> atom_serial;  // per atom
  std::vector<uint32_t> ring_rc;     // per ring -> resconfs
  int32_t ring_name = -1;            // id of "Ring"

  std::vector<std::string> strings;
  std::unordered_map<std::string, int32_t> string_ids;
  std::vector<std::string> warnings;

  int32_t intern(const std::string& s);
  void finish();
};

void parse_groups(const std::set<std::string>& all, const std::string& groups, std::set<std::string>& lig, std::set<std::string>& rec);
void pack_records(const arp_atom_record* recs, int64_t n, const std::string& groups, bool ignore_zero_occupancy, uint32_t flags,
                  Packed& out);

// rows (structure-local indices) -> sorted table; table.cpp
struct TableStore;
arph_table* build_table(const Packed& p, const arp_row* rows, int64_t n_rows);

// the integer side tables of arp_stage_frame (include/arpeggia_cuda.h) and the table that wraps a device-built frame
struct FrameTables {
  arp_frame_tables t{};
  std::vector<arp_frame_side> sides;
  std::vector<double> side_sc;
  std::vector<uint32_t> ent_side, ent_rank, ent_tie;
  std::vector<int32_t> ent_atomn, ent_atomi;
};
void build_frame_tables(const Packed& p, FrameTables& out);
// takes ownership of f (released by arph_table_free); t = the frame tables of p if the caller has them (a compact frame's
// per-entity columns are gathered through them; built here when NULL)
arph_table* table_from_frame(const Packed& p, arp_frame* f, const arp_frame_tables* t = nullptr);

}  // namespace arp

struct arph_packed;
const arp::Packed& arph_packed_ref(const arph_packed* p);
const arp_frame_tables* arph_packed_cached_frame_tables(const arph_packed* p);  // NULL until arph_packed_frame_tables built them
