/*
 * arpeggia_cuda.h -- C ABI of the B200 (sm_100a) contacts hot path: the drop-in boundary.
 *
 * This is what a Rust `arpeggia-cuda-sys` FFI crate (or cffi / ctypes) binds.  It replaces the three
 * search loops of the reference's `Interactions` trait and everything they call:
 *
 *   arp_contacts()  <->  InteractionComplex::get_atomic_contacts   src/contacts/complex.rs:189-299
 *                        InteractionComplex::get_ring_atom_contacts src/contacts/complex.rs:301-352
 *                        InteractionComplex::get_ring_ring_contacts src/contacts/complex.rs:354-405
 *                        (+ should_compare_entities/_residues        src/contacts/complex.rs:76-131,
 *                           find_vdw_contact/is_disulfide            src/contacts/vdw.rs:19-80,
 *                           find_ionic_bond/_repulsion               src/contacts/ionic.rs:11-35,
 *                           find_hydrogen_bond/find_weak_hydrogen_bond src/contacts/hbond.rs:30-110,
 *                           find_hydrophobic_contact                 src/contacts/hydrophobic.rs:10-24,
 *                           find_cation_pi / find_pi_pi              src/contacts/aromatic.rs:14-64,
 *                           pdbtbx create_hierarchy_rtree + rstar locate_within_distance)
 *
 * Plain pointers and sizes only; no C++/torch types.  Caller owns inputs (host memory, read-only for the
 * duration of a call); the library owns outputs until arp_result_free / the next staged call.  Every
 * function returns an arp_status; arp_last_error() gives the message.  Nothing throws or aborts across
 * the ABI.  A context is NOT thread-safe: one context per (device, host thread).  There is no CPU
 * fallback: without a CUDA device every compute entry point fails with ARP_ERR_CUDA.
 */
#ifndef ARPEGGIA_CUDA_H
#define ARPEGGIA_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum arp_status {
  ARP_OK = 0,
  ARP_ERR_IO = 1,
  ARP_ERR_BAD_GROUPS = 2,   /* utils.rs:77-80  "Invalid chain groups format! Use '/' for all-to-all comparisons." */
  ARP_ERR_EMPTY_GROUP = 3,  /* utils.rs:109-112 "Empty chain groups!" */
  ARP_ERR_NO_RINGS = 4,     /* complex.rs:50    "Error building ring positions" */
  ARP_ERR_ELEMENT = 5,      /* vdw.rs:24-28 unwrap on element / van_der_waals */
  ARP_ERR_NARROWING = 6,    /* mod.rs:142-178 u32 / i32 narrowing panics */
  ARP_ERR_RESIDUE_NAME = 7, /* residues.rs:132 unwrap on Residue::name() */
  ARP_ERR_ARG = 8,
  ARP_ERR_CUDA = 9,
  ARP_ERR_OOM = 10,
  ARP_ERR_LIMIT = 11,       /* > 2^31 atoms, > 2^32 cells, > 256 models, > 2^24 chains */
  ARP_ERR_NCCL = 12         /* NCCL not loadable, or a communicator call failed */
} arp_status;

/* Interaction labels: declaration order of `enum Interaction`, src/contacts/structs.rs:6-51. */
typedef enum arp_label {
  ARP_StericClash = 0, ARP_CovalentBond, ARP_Disulfide, ARP_VanDerWaalsContact, ARP_IonicBond,
  ARP_HydrogenBond, ARP_WeakHydrogenBond, ARP_PolarContact, ARP_WeakPolarContact, ARP_IonicRepulsion,
  ARP_SaltBridge, ARP_PiDisplacedStacking, ARP_PiTStacking, ARP_PiSandwichStacking,
  ARP_PiParallelInPlaneStacking, ARP_PiTiltedStacking, ARP_PiLStacking, ARP_CationPi,
  ARP_HydrophobicContact, ARP_N_LABELS
} arp_label;
const char* arp_label_name(int label); /* Display of Interaction, structs.rs:151-157 */

/* ---- per-atom attribute word (arp_structure.attr) ------------------------------------------------ */
#define ARP_ATTR_ELEM_MASK   0x0000000Fu /* element class, see arp_element_class(); 0 = hydrogen */
#define ARP_ATTR_LIGAND      0x00000010u /* chain is in the ligand set   (utils.rs:71-115) */
#define ARP_ATTR_RECEPTOR    0x00000020u /* chain is in the receptor set */
#define ARP_ATTR_POS_CONF    0x00000040u /* is_pos_ionizable(conformer name, atom name)  ionic.rs:83-91 */
#define ARP_ATTR_NEG_CONF    0x00000080u /* is_neg_ionizable(conformer name, atom name)  ionic.rs:94-99 */
#define ARP_ATTR_POS_RES     0x00000100u /* is_pos_ionizable(residue name, atom name)    aromatic.rs:18 */
#define ARP_ATTR_DONOR       0x00000200u /* is_hydrogen_donor    hbond.rs:162-181 */
#define ARP_ATTR_ACCEPTOR    0x00000400u /* is_hydrogen_acceptor hbond.rs:137-159 */
#define ARP_ATTR_WEAK_DONOR  0x00000800u /* element C and name != "C"  hbond.rs:204-207 */
#define ARP_ATTR_HYDROPHOBIC 0x00001000u /* is_hydrophobic(residue name, atom name) hydrophobic.rs:27-45 */
#define ARP_ATTR_CYS_SG      0x00002000u /* residue CYS, atom SG   vdw.rs:50-53 */
#define ARP_ATTR_GHOST       0x00004000u /* halo copy owned by another slab/rank: takes part in every test but never
                                            starts a row (the owner of the `from` atom / ring emits it) */

/* Element class for an upper-case element symbol, or -1 when pdbtbx would have no van der Waals radius
 * for it (the reference panics, vdw.rs:28).  Class 0 is hydrogen. */
int arp_element_class(const char* symbol);
/* covalent-single / van-der-Waals radius of a class (pdbtbx Element::atomic_radius) */
int arp_element_radii(int element_class, double* covalent_single, double* van_der_waals);

#define ARP_RING_FLAG 0x80000000u /* arp_row.from / .to: bit 31 set = ring index, else atom index */

/* One structure as SoA.  Atom arrays have n_atoms entries in atoms_with_hierarchy() order.  The pointers may be
 * host memory (pageable or page-locked) or device memory of the context's GPU (unified addressing decides the
 * copy direction), so buffers that were produced on the GPU -- e.g. a slab plus its NCCL-exchanged halo -- are
 * consumed in place. */
typedef struct arp_structure {
  int64_t n_atoms, n_residues, n_rings, n_h;
  const double* x;               /* angstrom, f64 exactly as parsed */
  const double* y;
  const double* z;
  const uint32_t* attr;          /* ARP_ATTR_* */
  const uint32_t* chain_key;     /* (model index << 24) | lexicographic rank of the chain id  (complex.rs:96,108,129) */
  const uint32_t* res_ordinal;   /* position of the residue in its chain  (build_residue_index, complex.rs:411-440) */
  const uint32_t* res_id;        /* index into the residue tables below */
  const uint32_t* res_h_begin;   /* n_residues+1: CSR into h_* = every hydrogen of the residue (hbond.rs:38-42) */
  const double* h_x;
  const double* h_y;
  const double* h_z;
  const int32_t* res_cb;         /* n_residues: atom index of the first "CB" / first "SG" of the residue, or -1 */
  const int32_t* res_sg;         /*             (is_disulfide, vdw.rs:55-74) */
  const double* ring_center;     /* 3*n_rings, xyz interleaved  (build_ring_positions, complex.rs:442-484) */
  const double* ring_normal;     /* 3*n_rings */
  const uint32_t* ring_attr;     /* n_rings: ARP_ATTR_LIGAND / ARP_ATTR_RECEPTOR of the ring's chain */
  const uint32_t* ring_chain_key;
  const uint32_t* ring_res_ordinal;
  /* Optional hint: the axis-aligned bounding box of x / y / z, computed by whoever packed the atoms (the host packer
   * walks every atom anyway).  With has_bbox != 0 the cell grid is laid out without a device reduction and a host round
   * trip.  The box is verified on the device: if an atom lies outside it the job silently falls back to the computed
   * box.  Leave has_bbox = 0 for device-resident or unpacked coordinates. */
  double bbox_min[3];
  double bbox_max[3];
  int32_t has_bbox;
  int32_t reserved_;
  /* Optional compact form: chain_key and res_ordinal are the same for every atom of a residue.  When BOTH per-residue
   * arrays are given (n_residues entries, indexed by res_id), the per-atom chain_key / res_ordinal arrays are not read
   * (they may be NULL) and the device expands the values while it sorts the atoms: 8 bytes per atom less to move over
   * PCIe, a fifth of the upload. */
  const uint32_t* res_chain_key;
  const uint32_t* res_res_ordinal;
  /* Optional: the atoms of residue r are [res_atom_begin[r], res_atom_begin[r+1]) (n_residues+1 entries; true for
   * atoms_with_hierarchy() order).  When given, res_id is not read (may be NULL): the device expands it -- another
   * 4 bytes per atom that stay off PCIe. */
  const uint32_t* res_atom_begin;
  /* Optional on-lattice form of the atom coordinates: integer thousandths of an angstrom, the resolution of a PDB
   * fixed-column field (columns 31-54, "%8.3f").  The coordinate is (double)v / 1000.0 with one IEEE division -- the
   * correctly rounded double of the decimal v/1000, i.e. the value a correctly rounded parser returns for that field.
   * When ALL THREE are given, x / y / z are not read (may be NULL) and the device rebuilds the f64 coordinates itself:
   * 12 instead of 24 bytes per atom over PCIe.  The packer only offers this form after checking every coordinate of
   * the structure against that identity bit for bit; hydrogens (h_*) and ring planes stay f64. */
  const int32_t* x_milli;
  const int32_t* y_milli;
  const int32_t* z_milli;
  /* Optional ids of the atoms / rings in a larger, undivided structure (n_atoms / n_rings entries, < 2^31).  When
   * given, arp_row.from / .to carry these ids instead of job-local indices.  Required for slab jobs (arp_slab_exchange):
   * halo copies arrive with the ids their owner gave them, so the union of the ranks' rows is the row set of the
   * undivided structure. */
  const uint32_t* atom_gid;
  const uint32_t* ring_gid;
  /* Optional per-residue tag (n_residues entries) with the property
   *     |res_tag[a] - res_tag[b]| <= 1   <=>   a and b are in the same model and chain and their residue ordinals
   *                                            differ by at most 1                          (complex.rs:108-120)
   * e.g. (index of the residue in atoms_with_hierarchy() order) + 2 * (index of its chain).  The emit-radius screen
   * decides the in-chain neighbour rule from it without gathering chain keys.  When NULL the device derives the tags
   * from res_chain_key / res_res_ordinal, ASSUMING atoms_with_hierarchy() order: the residues of a chain are contiguous
   * and in ordinal order (violations that are visible between neighbouring residues fail the job with ARP_ERR_ARG).
   * Slab jobs (arp_slab_exchange) need tags that are valid across ranks, so there the caller must supply them. */
  const uint32_t* res_tag;
  /* Optional most compact form of the two per-atom arrays that are left (single-structure jobs; what a host keeps when
   * the PCIe copy is the cost of a call):
   *   xyz_milli21  one 64-bit word per atom:  dx | dy << 21 | dz << 42,  d* = *_milli - xyz_milli21_origin[*] in [0, 2^21)
   *                -- the on-lattice coordinates of a structure up to 2 097 A across, 8 instead of 12 bytes per atom;
   *                when given, x_milli / y_milli / z_milli and x / y / z are not read (may be NULL);
   *   attr16       attr as 16-bit words (every ARP_ATTR_* bit a caller sets lies below bit 15), 2 instead of 4 bytes;
   *                when given, attr is not read (may be NULL).
   * The device rebuilds x / y / z (one IEEE division per coordinate, as for x_milli) and attr itself. */
  const uint64_t* xyz_milli21;
  int32_t xyz_milli21_origin[3];
  int32_t reserved2_;
  const uint16_t* attr16;
} arp_structure;

#define ARP_FLAG_EMIT_RADIUS_SCREEN 0x1u /* screen atom pairs at min(dist_cutoff, largest distance at which any
                                            label can fire) instead of dist_cutoff. Rows are identical;
                                            pairs_in_cutoff is then not computed (reported as -1). */
#define ARP_FLAG_COUNT_CLASSIFIED    0x2u /* also count pairs_classified: every pair within dist_cutoff is sent through
                                            the exact filter, not only those close enough to produce a row (slower;
                                            without it pairs_classified is reported as -1). Rows are identical. */
typedef struct arp_params {
  double vdw_comp;     /* -c / vdw_comp   (cli/contacts.rs:37-39) */
  double dist_cutoff;  /* -d / dist_cutoff (cli/contacts.rs:41-43) */
  uint32_t flags;
  uint32_t reserved_;
} arp_params;

/* One output row = one ResultEntry (structs.rs:73-85) before string decoding. 16 bytes. */
typedef struct arp_row {
  uint32_t from;       /* atom index within the job (batch: global over the concatenated structures) */
  uint32_t to;
  uint32_t label;      /* arp_label */
  float distance;      /* (float) of the f64 distance, round-to-nearest-even == Rust `as f32` (mod.rs:148) */
} arp_row;

typedef struct arp_result {
  int64_t n_rows;
  const arp_row* rows;            /* page-locked host memory owned by THIS result (valid until arp_result_free, whatever
                                     else is called on the context meanwhile); UNSORTED (order is not deterministic) */
  int64_t n_structures;
  const int64_t* atom_begin;      /* n_structures+1: offsets of each structure's atoms in the job */
  const int64_t* ring_begin;      /* n_structures+1 */
  int64_t pairs_screened;         /* distance tests evaluated by the neighbour kernel (unordered candidate pairs) */
  int64_t pairs_in_cutoff;        /* ordered pairs (from in ligand, non-H; to != from) with d^2 <= rc^2, or -1 */
  int64_t pairs_classified;       /* ordered pairs that passed should_compare_entities, or -1 (see flags) */
} arp_result;

/* Device-side timings and sizes of the last job (CUDA events on the context's stream). */
typedef struct arp_stats {
  int64_t n_atoms, n_rings, n_cells, n_nonempty_cells, n_rows;
  int64_t kernel_launches;        /* kernels of this library launched by the last run */
  int32_t pair_kernel_retries;    /* re-runs because the row buffer or a pair queue was too small */
  int32_t upload_copies;          /* host->device copies the last upload issued for the job's arrays (neighbours in host memory travel as one) */
  int64_t n_screened_pairs;       /* entries of the first pair queue (k_screen -> k_filter), unused window tails included */
  int64_t n_filtered_pairs;       /* entries of the second pair queue (k_filter -> k_exact) */
  float ms_h2d, ms_grid, ms_pairs, ms_rings, ms_d2h, ms_run; /* ms_run = the whole job; ms_rings overlaps ms_pairs (side stream) */
  float ms_screen, ms_filter, ms_exact, reserved2_;          /* the three kernels of ms_pairs */
  /* host wall clock of the phases of the last one-shot call (arp_contacts / arp_contacts_batch), each from its first API
   * call to the return of its last synchronisation: upload (enqueueing the copies: a one-shot call does not wait for
   * them), run (enqueue + wait for copies, kernels and counters), download.  Their sum minus ms_h2d + ms_run + ms_d2h
   * is host overhead (submission, wake-ups after a synchronisation). */
  float ms_wall_upload, ms_wall_run, ms_wall_download, ms_wall_call;
} arp_stats;

typedef struct arp_ctx arp_ctx;

int  arp_ctx_create(arp_ctx** out, int device_ordinal);
void arp_ctx_destroy(arp_ctx* ctx);
const char* arp_last_error(const arp_ctx* ctx); /* ctx may be NULL: error of the last failed create */
void* arp_ctx_stream(arp_ctx* ctx);             /* the cudaStream_t every kernel is launched on (for event timing) */

/* One-shot: host SoA in, host rows out (H2D, all kernels, D2H, stream-synchronised before return). */
int  arp_contacts(arp_ctx* ctx, const arp_structure* s, const arp_params* p, arp_result** out);
/* Many independent structures in one job; no pair ever spans two structures. */
int  arp_contacts_batch(arp_ctx* ctx, const arp_structure* const* s, int64_t n, const arp_params* p, arp_result** out);
void arp_result_free(arp_ctx* ctx, arp_result* r); /* ctx: the (live) context the result came from -- its row buffer is kept
                                                       for that context's next download -- or NULL (e.g. the context is
                                                       already destroyed): the buffer is released */

/* Staged form of the same job, for pipelines that keep inputs resident in HBM:
 * upload once, run any number of times (asynchronous on the stream), download the last run's rows. */
int  arp_stage_upload(arp_ctx* ctx, const arp_structure* const* s, int64_t n);
int  arp_stage_run(arp_ctx* ctx, const arp_params* p);
int  arp_stage_download(arp_ctx* ctx, arp_result** out);
int  arp_stage_stats(arp_ctx* ctx, arp_stats* out); /* synchronises the stream */

/* ---- weighted neighbour sum over the resident atoms -----------------------------------------------
 *   arp_stage_neighbor_sum()  <->  the radius query of get_per_atom_sap_score            src/sap.rs:181-215
 * After arp_stage_upload (no run needed): for every atom x with select[x] != 0
 *     out[x] = sum of weight[y] over the atoms y OF THE SAME STRUCTURE with select[y] != 0 and
 *              distance_2(x, y) <= (double)(radius * radius),
 * x itself included (sap.rs:193-194), the comparison inclusive (rstar locate_within_distance) and the squared radius
 * formed in f32 like sap.rs:184; out[x] = 0 where select[x] == 0.  weight / select / out have n_atoms entries in the
 * job's atom order (a batch job: the structures' atoms one after the other) and may be host or device memory.  In the SAP score select = h.is_sidechain() and
 * weight = hydrophobicity(resn) * clamp(sasa / max_asa(resn), 0, 1) (sap.rs:199-210): computing the SASA is outside this
 * library.  The distance test is exact FP64 in the reference's operation order, so the neighbour SET is the reference's;
 * the sum is accumulated in f64 and rounded once (the reference adds f32 terms in the R*-tree's unspecified iteration
 * order: equal for exactly representable sums, within n * 2^-24 * sum|w| otherwise). */
int  arp_stage_neighbor_sum(arp_ctx* ctx, const float* weight, const unsigned char* select, float radius, float* out);
/* device times of the last arp_stage_neighbor_sum: cell-list build, sum kernel; cells of its grid */
int  arp_neighbor_sum_stats(arp_ctx* ctx, float* ms_grid, float* ms_sum, int64_t* n_cells);

/* ---- the 20-column frame, built on the device -----------------------------------------------
 *   arp_stage_frame()  <->  results_to_df                          src/contacts/mod.rs:140-181
 *                           collect_sc_stats + sc_results_to_df    src/contacts/complex.rs:137-174, mod.rs:184-214
 *                           concat + left join + 10-key sort       src/contacts/mod.rs:111-136
 * The rows of the last run never leave the GPU: they are keyed (one 128-bit key per row from dense integer ranks
 * the caller prepares once per structure), radix-sorted, decoded into dictionary-encoded columns and joined with the
 * side-chain plane statistics on the device; only the finished columns cross PCIe.  String columns are int32 ids
 * into the CALLER's dictionary (the library never sees a string).  Single-structure jobs only.
 *
 * An "entity" is what a row end points at: atoms 0..n_atoms-1, then rings n_atoms..n_atoms+n_rings-1
 * (InteractingEntity, structs.rs:54-71; rings carry atomn "Ring", atomi 0, complex.rs:334-342).
 * A "side" is one (model, chain, residue, conformer): the fields of InteractingEntity that are not per atom. */
typedef struct arp_frame_side { /* 32 bytes */
  int32_t  chain, resn, resi, icode, altloc; /* string ids; resi already narrowed to i32 (mod.rs:150-178) */
  uint32_t model;      /* model serial, narrowed to u32 (mod.rs:142-146) */
  uint32_t chain_rank; /* dense rank of (model serial, chain id bytes): sort keys model, from_chain, to_chain */
  uint32_t has_sc;     /* a side-chain plane exists for this residue (complex.rs:486-514) */
} arp_frame_side;

#define ARP_ENT_NARROW 0x80000000u /* ent_tie bit 31: a field of this entity does not fit the frame's integer types;
                                      a row that references it fails the call with ARP_ERR_NARROWING (the reference panics) */
typedef struct arp_frame_tables {
  int64_t n_atoms, n_rings, n_sides;
  const arp_frame_side* sides; /* [n_sides] */
  const double*   side_sc;     /* [6 * n_sides] side-chain plane: centre xyz, normal xyz (ignored where !has_sc) */
  const uint32_t* ent_side;    /* [n_atoms + n_rings] entity -> side */
  const int32_t*  ent_atomn;   /* string id of the atom name */
  const int32_t*  ent_atomi;   /* atom serial number as i32; 0 for rings */
  const uint32_t* ent_rank;    /* dense rank of (resi, altloc bytes, atomi) among the entities of the same chain_rank:
                                  sort keys 4-6 and 7-9 (only compared once the chain keys are equal) */
  const uint32_t* ent_tie;     /* position among the entities of equal (chain_rank, ent_rank), in entity order (makes
                                  the order total; the reference leaves such ties unspecified) | ARP_ENT_NARROW */
  uint8_t  label_rank[32];     /* rank of arp_label_name(l) in byte order: sort key 10 (the column is a string) */
  uint32_t n_chain_ranks, n_ent_ranks, max_tie; /* exclusive upper bounds of chain_rank / ent_rank, largest ent_tie */
} arp_frame_tables;

/* Library-owned, page-locked host memory; valid until arp_frame_free.  Rows are in the frame's final order. */
typedef struct arp_frame {
  int64_t   n_rows;
  uint32_t* model;
  uint8_t*  interaction; /* arp_label */
  float*    distance;
  int32_t  *from_chain, *from_resn, *from_resi, *from_insertion, *from_altloc, *from_atomn, *from_atomi;
  int32_t  *to_chain,   *to_resn,   *to_resi,   *to_insertion,   *to_altloc,   *to_atomn,   *to_atomi;
  float    *sc_centroid_dist, *sc_dihedral, *sc_centroid_angle;
  uint8_t  *sc_valid; /* 0 = null (either residue has no plane, mod.rs:112-122); 1 = values present; 2 = present, but
                         an angle sits within 4 ulp (f64) of an f32 rounding boundary or is NaN: a caller that
                         needs the host libm's bits recomputes these rows (arph_contacts does) */
  uint32_t *from, *to; /* the row ends as arp_row carries them (entity = v & ARP_RING_FLAG ? n_atoms + index : v) */
  int64_t pairs_in_cutoff, pairs_classified;
  float ms_sort, ms_columns, ms_d2h; /* device times of the key + radix-sort passes, the column kernel, the copy */
  int32_t sort_passes, key_bits;
} arp_frame;

/* After arp_stage_run of a single-structure job: sort + decode + join on the device, columns to the host. */
int  arp_stage_frame(arp_ctx* ctx, const arp_frame_tables* t, arp_frame** out);
/* One-shot: arp_stage_upload + arp_stage_run + arp_stage_frame. */
int  arp_contacts_frame(arp_ctx* ctx, const arp_structure* s, const arp_params* p, const arp_frame_tables* t, arp_frame** out);
/* The compact frame: the same sort, the same rows in the same order, but the fifteen per-entity columns (`model` and the
 * fourteen from_ / to_ columns) are NULL and never leave the device -- they are pure gathers through `from` / `to`
 * (entity -> side -> field) the host can do from the tables it built (arph_table_from_frame does, on the host's threads).
 * What crosses PCIe is from, to, interaction, distance, the three sc_* and sc_valid: 26 bytes per row instead of 86;
 * t->ent_atomn / t->ent_atomi may be NULL and are not uploaded.  Errors (narrowing, rows outside the tables) as above. */
int  arp_stage_frame_compact(arp_ctx* ctx, const arp_frame_tables* t, arp_frame** out);
int  arp_contacts_frame_compact(arp_ctx* ctx, const arp_structure* s, const arp_params* p, const arp_frame_tables* t, arp_frame** out);
void arp_frame_free(arp_ctx* ctx, arp_frame* f); /* ctx may be NULL; the memory returns to a process-wide pool */

/* ---- one assembly cut into x-slabs over several GPUs (BASELINE config 5) -------------------------------
 * Partitions the three search loops of complex.rs:197-213, 306-321, 356-370 by the owner of the row's `from` entity.
 * Every rank (one context = one GPU) owns a set of WHOLE residues -- normally those of an x-slab -- and uploads them with
 * arp_stage_upload (one structure, compact per-residue forms, atom_gid / ring_gid / res_tag given, chain keys and residue
 * ordinals those of the undivided structure).  arp_slab_exchange then completes the job with read-only halo copies:
 *   k_halo_mark      every owned residue with an atom whose x lies within `width` of a peer's owned x-range (and every
 *                    ring whose centroid does) is marked for that peer; width = max(|dist_cutoff|, 6.0) (pi-pi reaches
 *                    6.0 A whatever the cutoff, aromatic.rs:33-64) -- whole residues, so the hydrogen lists and CB / SG
 *                    slots the per-pair tests gather from (hbond.rs:38-42, vdw.rs:55-74) are complete on the receiver
 *   k_halo_count/_offsets   deterministic message offsets (block sums + one small scan), counts to the host (the one
 *                    host synchronisation of the exchange: NCCL needs the byte counts)
 *   k_halo_pack      residues, atoms, hydrogens, rings -> one contiguous message per peer
 *   ncclSend/ncclRecv (grouped, on the context's stream; NVLink 5 / NVSwitch) -- or peer copies, see below
 *   k_halo_unpack    appends the received records to the resident job, marks them ARP_ATTR_GHOST (they take part in
 *                    every test but never start a row) and rebases their indices
 * after which arp_stage_run / arp_stage_download work unchanged; rows carry global ids.  Ghosts of an earlier exchange
 * are dropped first, so the call can be repeated.  There is no torch or host-side selection anywhere on this path.
 *
 * Communicators: one process per GPU -> arp_comm_get_id on one rank, the 128 bytes distributed by the host's own means
 * (MPI, a socket, torch.distributed), arp_ctx_comm_init on every rank.  One process driving several GPUs ->
 * arp_ctx_create_multi (ncclCommInitAll) and arp_slab_exchange_multi.  NCCL is loaded with dlopen("libnccl.so.2") on
 * first use (the copy already in the process, if any), so the library itself has no NCCL link dependency.
 * arp_ctx_create_multi accepts a device more than once (NCCL does not): such a set moves its messages with
 * cudaMemcpyPeerAsync instead -- the same kernels and byte layout, used to test the path on a single GPU. */
typedef struct arp_comm_id { char bytes[128]; } arp_comm_id; /* ncclUniqueId */
int  arp_comm_get_id(arp_comm_id* out);
int  arp_ctx_comm_init(arp_ctx* ctx, const arp_comm_id* id, int rank, int world);
int  arp_ctx_create_multi(arp_ctx** out /* [n_devices] */, const int* devices, int n_devices);

typedef struct arp_slab_info {
  int64_t own_atoms, own_residues, own_rings;
  int64_t ghost_atoms, ghost_residues, ghost_rings, ghost_h;
  int64_t sent_atoms, sent_residues, sent_rings;
  int64_t bytes_sent, bytes_received;
  int32_t n_peers;       /* ranks this one exchanged a (possibly empty) message with */
  int32_t transport;     /* 1 = NCCL send/recv, 2 = peer copies (duplicate devices in one process) */
  float ms_select, ms_transfer, ms_unpack, ms_total; /* CUDA events on the context's stream; ms_total includes the host wait */
} arp_slab_info;
int  arp_slab_exchange(arp_ctx* ctx, const arp_params* p, arp_slab_info* info /* may be NULL */);
int  arp_slab_exchange_multi(arp_ctx* const* ctx, int n, const arp_params* p, arp_slab_info* info /* [n] or NULL */);

/* ---- the neighbour sum for callers that cannot allocate -----------------------------------------------------------------
 * No counterpart in the reference: added for callers that must not call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same computation as
 * arp_stage_neighbor_sum -- the same kernels -- on caller-owned DEVICE arrays of one structure:
 *     out[i] = select[i] ? sum of weight[j] over j with select[j] and distance_2(i, j) <= (double)(radius * radius) : 0
 * with every byte of scratch (bounding box, cell list, scan status, sorted records) in one file-scope __device__ array
 * of fixed size: ARP_STATIC_MAX_ATOMS atoms at most (more fails with ARP_ERR_LIMIT; the cell grid is coarsened to fit
 * ARP_STATIC_MAX_CELLS).  Every GPU that loads the library holds that array (8.4 MB) until the process exits.
 * The call allocates nothing -- no cudaMalloc / cudaFree / cudaMallocHost, no context, no library routine that does --
 * launches on `stream` (a cudaStream_t; NULL = the default stream) of `device` and returns after synchronising that
 * stream.  x / y / z / weight / select / out are device pointers of n_atoms entries.  ALL calls on a device share the
 * one array (it is re-initialised by every call): calls that overlap on the GPU -- from two host threads or two
 * streams -- race and give wrong sums; serialise them.  Errors: arp_last_error(NULL). */
#define ARP_STATIC_MAX_ATOMS 131072
#define ARP_STATIC_MAX_CELLS 262144
int  arp_neighbor_sum_static(int device, void* stream, const double* x, const double* y, const double* z, const float* weight,
                             const unsigned char* select, int64_t n_atoms, float radius, float* out);

/* Page-locked host memory for callers that want true asynchronous H2D/D2H (optional). */
int  arp_host_alloc(void** out, int64_t bytes);
void arp_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif
