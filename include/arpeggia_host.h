/*
 * arpeggia_host.h -- C ABI of the host side of the contacts path: what sits either side of arp_contacts().
 *
 * Mirrors the reference's host functions for this path (same argument meaning and error behaviour):
 *
 *   arph_parse_pdb / _cif / _file <-> pdbtbx ReadOptions{only_atomic_coords, Loose}.read()   src/utils.rs:53-57
 *   arph_pack             <->  load_model residue filter                              src/utils.rs:60
 *                              optional zero-occupancy filter                         src/python.rs:45-47
 *                              InteractionComplex::new (parse_groups, build_residue_index,
 *                              build_ring_positions, build_sc_plane_positions)        src/contacts/complex.rs:36-68, 411-514
 *   arph_build_table      <->  results_to_df, collect_sc_stats, sc_results_to_df,
 *                              concat + left join + 10-key sort                      src/contacts/mod.rs:79-214, complex.rs:137-174
 *   arph_contacts[_file]  <->  get_contacts / arpeggia.contacts()                     src/contacts/mod.rs:61, src/python.rs:31-56
 *
 * The atom-record and table layouts are byte-identical to the oracle's (oracle/oracle.h) on purpose, so a
 * test can hand the same buffers to both; the two code bases share no source.
 */
#ifndef ARPEGGIA_HOST_H
#define ARPEGGIA_HOST_H

#include <stdint.h>

#include "arpeggia_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* One ATOM/HETATM record as a parser yields it. 96 bytes. */
typedef struct arp_atom_record {
  double  x, y, z;
  double  occupancy;
  int64_t serial;        /* atom serial number */
  int64_t resi;          /* residue sequence number */
  int64_t model;         /* model serial number; 0 when the file has no MODEL record */
  char    chain[8];      /* NUL padded */
  char    resn[8];
  char    name[8];
  char    element[4];    /* upper-case symbol, "" = unknown */
  char    altloc[2];
  char    icode[2];
  uint8_t hetero;
  uint8_t pad_[7];
} arp_atom_record;

/* The 20-column frame of get_contacts (mod.rs:111-136).  String columns are int32 indices into `strings`. */
typedef struct arph_table {
  int64_t n_rows;
  uint32_t* model;
  uint8_t*  interaction;           /* arp_label; the column value is arp_label_name(code) */
  float*    distance;
  int32_t  *from_chain, *from_resn, *from_resi, *from_insertion, *from_altloc, *from_atomn, *from_atomi;
  int32_t  *to_chain,   *to_resn,   *to_resi,   *to_insertion,   *to_altloc,   *to_atomn,   *to_atomi;
  float    *sc_centroid_dist, *sc_dihedral, *sc_centroid_angle;
  uint8_t  *sc_valid;              /* 0 = the three sc_* values are null (left join miss, mod.rs:112-122) */
  int32_t n_strings;
  const char* const* strings;
  int64_t n_atoms;
  int64_t pairs_in_cutoff;
  int64_t pairs_classified;
  double  t_setup_s, t_search_s, t_classify_s, t_rings_s, t_table_s; /* pack, -, device run, -, table */
  const uint8_t* sc_validity;      /* the same as sc_valid != 0, as a bitmap (bit k & 7 of byte k >> 3: Arrow's layout) */
  int64_t sc_null_count;           /* rows with sc_valid == 0 */
} arph_table;

#define ARPH_FLAG_NO_MULTIMODEL_QUIRK 1u /* do NOT reproduce complex.rs:447-449 (planes of every model = last model's) */
#define ARPH_FLAG_SKIP_LOAD_FILTER    2u /* records are already restricted to the 20 amino acids + HOH */
#define ARPH_FLAG_NO_COUNTERS          4u /* arph_contacts: the caller only wants the frame (what get_contacts returns).  The
                                            device then screens at the emit radius (ARP_FLAG_EMIT_RADIUS_SCREEN) and the
                                            table's pairs_in_cutoff / pairs_classified are -1.  Rows are identical. */

#define ARPH_FLAG_DEVICE_COLUMNS        8u /* arph_contacts: decode the fifteen per-entity columns on the device
                                            (arp_contacts_frame, 86 bytes per row over PCIe) instead of gathering them on
                                            the host from a compact frame (arp_contacts_frame_compact, 26 bytes per row).
                                            The table is the same either way. */

const char* arph_last_error(void);

/* Fixed-column PDB reader (ATOM / HETATM / MODEL). *out is malloc'd; release with arph_free. */
int  arph_parse_pdb(const char* path, arp_atom_record** out, int64_t* n);
/* mmCIF: the _atom_site loop.  Chain ids and residue numbers come from the LABEL identifiers (label_asym_id,
 * label_seq_id; auth_* only where the label value is absent or null) -- pdbtbx 0.12's column choice as far as it can be
 * established without its source (not under /root/reference: parity unpinned).  arph_parse_cif_ids with
 * use_auth_ids != 0 takes the author identifiers instead (what the PDB-format file of the same entry carries). */
int  arph_parse_cif(const char* path, arp_atom_record** out, int64_t* n);
int  arph_parse_cif_ids(const char* path, int use_auth_ids, arp_atom_record** out, int64_t* n);
/* Dispatch on the extension like pdbtbx (src/utils.rs:51-57): .cif / .mmcif -> mmCIF, anything else -> PDB. */
int  arph_parse_file(const char* path, arp_atom_record** out, int64_t* n);
void arph_free(void* p);

/* Packed structure: hierarchy + device SoA + the dictionaries needed to decode rows. */
typedef struct arph_packed arph_packed;
int  arph_pack(const arp_atom_record* atoms, int64_t n, const char* groups, int ignore_zero_occupancy,
               uint32_t flags, arph_packed** out);
/* count independent structures, one per host thread (BASELINE config 4); out[count].  On failure every out[k] is NULL and
 * the message names the first structure that failed. */
int  arph_pack_many(const arp_atom_record* const* atoms, const int64_t* n, int64_t count, const char* groups,
                    int ignore_zero_occupancy, uint32_t flags, arph_packed** out);
const arp_structure* arph_packed_structure(const arph_packed* p);
int64_t arph_packed_n_warnings(const arph_packed* p);           /* ring residues with < 3 ring atoms (mod.rs:64-68) */
const char* arph_packed_warning(const arph_packed* p, int64_t i);
void arph_packed_free(arph_packed* p);
/* A packed structure as one PART of a larger one (one rank's slab of an assembly, arp_slab_exchange): the packer ranks
 * chain ids among the chains IT sees (complex.rs:129 compares the id strings), so a host that packs only its part
 * replaces those ranks by the ranks in the undivided structure, and attaches the atoms' / rings' ids there plus residue
 * tags that are valid across parts (arp_structure.atom_gid / ring_gid / res_tag; res_tag NULL keeps the local tags). */
int64_t arph_packed_n_chains(const arph_packed* p);
const char* arph_packed_chain_id(const arph_packed* p, int64_t rank);  /* chain ids in byte order = local rank order */
int  arph_packed_remap_chain_ranks(arph_packed* p, const uint32_t* new_rank, int64_t n_chains);
int  arph_packed_set_global_ids(arph_packed* p, const uint32_t* atom_gid, const uint32_t* ring_gid, const uint32_t* res_tag);
/* The integer side tables arp_stage_frame / arp_contacts_frame need (dense sort ranks of the string keys, per-entity
 * and per-residue columns).  Built on first use, owned by the packed structure, valid until arph_packed_free. */
int  arph_packed_frame_tables(arph_packed* p, const arp_frame_tables** out);
/* Wrap a device-built frame as the host table (no column is copied; takes ownership of f, also on failure).  A compact
 * frame's model / from_* / to_* columns are gathered here (entity -> side -> field), rows split over the host's threads. */
int  arph_table_from_frame(const arph_packed* p, arp_frame* f, arph_table** out);

/* rows (structure-local atom / ring indices) -> sorted 20-column table. */
int  arph_build_table(const arph_packed* p, const arp_row* rows, int64_t n_rows, arph_table** out);
void arph_table_free(arph_table* t);

/* get_contacts: pack -> arp_contacts -> table. */
int  arph_contacts(arp_ctx* ctx, const arp_atom_record* atoms, int64_t n, const char* groups, double vdw_comp,
                   double dist_cutoff, int ignore_zero_occupancy, uint32_t flags, arph_table** out);
int  arph_contacts_file(arp_ctx* ctx, const char* path, const char* groups, double vdw_comp, double dist_cutoff,
                        int ignore_zero_occupancy, uint32_t flags, arph_table** out);

/* Plane helpers (residues.rs:31-75, 270-298), exposed for the known-answer tests. */
int    arph_plane_fit(const double* xyz, int64_t n_points, double center[3], double normal[3]);
double arph_plane_point_angle(const double center[3], const double normal[3], const double point[3]);
double arph_plane_dihedral(const double n1[3], const double n2[3]);
int    arph_parse_groups(const char* chains_csv, const char* groups, char* ligand_out, char* receptor_out, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif
