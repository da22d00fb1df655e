/*
 * oracle.h -- C interface of the CPU oracle for the arpeggia contacts hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker / the timed CPU baseline.
 *
 * The oracle is a plain C++17 / FP64 restatement of the reference's
 * src/contacts/ (see arpeggia_oracle.cpp for per-function file:line citations).
 * Parity pinning: see the header of arpeggia_oracle.cpp.
 *
 * The record and table layouts are deliberately byte-identical to the product's
 * (include/arpeggia_host.h) so one numpy dtype drives both, but this header is
 * standalone: the oracle includes no product header and links no product code.
 */
#ifndef ARPEGGIA_ORACLE_H
#define ARPEGGIA_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One ATOM/HETATM record as a parser yields it (pdbtbx `Atom` + its hierarchy ids). 96 bytes. */
typedef struct orc_atom_record {
  double  x, y, z;       /* angstrom */
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
} orc_atom_record;

/* Interaction codes: declaration order of `enum Interaction`, src/contacts/structs.rs:6-51. */
enum {
  ORC_StericClash = 0, ORC_CovalentBond, ORC_Disulfide, ORC_VanDerWaalsContact, ORC_IonicBond,
  ORC_HydrogenBond, ORC_WeakHydrogenBond, ORC_PolarContact, ORC_WeakPolarContact,
  ORC_IonicRepulsion, ORC_SaltBridge, ORC_PiDisplacedStacking, ORC_PiTStacking,
  ORC_PiSandwichStacking, ORC_PiParallelInPlaneStacking, ORC_PiTiltedStacking, ORC_PiLStacking,
  ORC_CationPi, ORC_HydrophobicContact, ORC_N_INTERACTIONS
};

/* The 20-column result of get_contacts (src/contacts/mod.rs:61-136), dictionary encoded.
 * String columns are int32 indices into the NUL-terminated string arrays below. */
typedef struct orc_table {
  int64_t n_rows;
  uint32_t* model;
  uint8_t*  interaction;           /* code above; name = orc_interaction_name(code) */
  float*    distance;
  int32_t  *from_chain, *from_resn, *from_resi, *from_insertion, *from_altloc, *from_atomn, *from_atomi;
  int32_t  *to_chain,   *to_resn,   *to_resi,   *to_insertion,   *to_altloc,   *to_atomn,   *to_atomi;
  float    *sc_centroid_dist, *sc_dihedral, *sc_centroid_angle;
  uint8_t  *sc_valid;              /* 0 = the three sc_* values are null */
  int32_t n_strings;               /* one shared dictionary for every string column */
  const char* const* strings;
  /* workload counters (for the metric) */
  int64_t n_atoms;                 /* atoms after load_model filter (all conformer copies) */
  int64_t pairs_in_cutoff;         /* ordered pairs, from-atom in ligand & non-H, any to-atom != itself, d2 <= rc^2 */
  int64_t pairs_classified;        /* ordered pairs that passed every filter */
  double  t_setup_s, t_search_s, t_classify_s, t_rings_s, t_table_s;
} orc_table;

const char* orc_last_error(void);
const char* orc_interaction_name(int code);

/* Minimal fixed-column PDB reader (ATOM/HETATM/MODEL/ENDMDL). *out is malloc'd; free with orc_free. */
int  orc_parse_pdb(const char* path, orc_atom_record** out, int64_t* n);
void orc_free(void* p);

/* flags */
#define ORC_FLAG_NO_MULTIMODEL_QUIRK 1u  /* do NOT reproduce complex.rs:447-449 (planes taken from the last model) */
#define ORC_FLAG_SKIP_LOAD_FILTER    2u  /* skip utils.rs:60 residue filter (records are already filtered) */
#define ORC_FLAG_NO_TABLE            4u  /* counters and timings only; n_rows is set, columns are NULL */

/* get_contacts on a list of atom records (= load_model + optional zero-occupancy filter + get_contacts).
 * num_threads: 1 = the reference default (-j 1); 0 = all cores. Returns 0 or an error code. */
int  orc_contacts(const orc_atom_record* atoms, int64_t n, const char* groups, double vdw_comp,
                  double dist_cutoff, int ignore_zero_occupancy, uint32_t flags, int num_threads,
                  orc_table** out);
void orc_table_free(orc_table* t);

/* Known-answer helpers. */
int  orc_plane_fit(const double* xyz, int64_t n_points, double center[3], double normal[3]);
double orc_plane_point_angle(const double center[3], const double normal[3], const double point[3]);
double orc_plane_point_dist(const double center[3], const double point[3]);
double orc_plane_dihedral(const double n1[3], const double n2[3]);
/* chains_csv: every chain id, comma separated. Writes sorted comma separated ligand / receptor ids. */
int  orc_parse_groups(const char* chains_csv, const char* groups, char* ligand_out, char* receptor_out, int64_t cap);
/* Residue planes after load_model: kind 0 = ring_atoms, 1 = sc_plane_atoms (residues.rs:163-298).
 * Looks up the first residue in file order with this chain id and resi. n_atoms_out = atoms used. */
int  orc_residue_plane(const orc_atom_record* atoms, int64_t n, const char* chain, int64_t resi, int kind,
                       int64_t* n_atoms_out, double center[3], double normal[3]);

/* The radius query of the SAP score, src/sap.rs:181-215, on caller-supplied weights: for every atom x with select[x]
 *   out[x] = sum of weight[y] over the atoms y with select[y] and distance_2(x, y) <= (double)(radius * radius)
 * (x included; rstar locate_within_distance is inclusive; sap.rs:184 squares the radius in f32), summed in f32 in
 * increasing atom order (the R*-tree's iteration order is unspecified); out[x] = 0 where !select[x].
 * In the reference select = is_sidechain() and weight = hydrophobicity * clamp(sasa / max_asa, 0, 1). */
int  orc_neighbor_sum(const double* x, const double* y, const double* z, int64_t n, const float* weight,
                      const unsigned char* select, float radius, float* out);

/* error codes */
enum {
  ORC_OK = 0, ORC_ERR_IO = 1, ORC_ERR_BAD_GROUPS = 2, ORC_ERR_EMPTY_GROUP = 3, ORC_ERR_NO_RINGS = 4,
  ORC_ERR_ELEMENT = 5, ORC_ERR_NARROWING = 6, ORC_ERR_RESIDUE_NAME = 7, ORC_ERR_ARG = 8
};

#ifdef __cplusplus
}
#endif
#endif
