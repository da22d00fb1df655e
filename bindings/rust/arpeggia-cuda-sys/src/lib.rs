//! Raw bindings to `include/arpeggia_cuda.h` — the device path that replaces the three search loops of
//! `impl Interactions for InteractionComplex` (src/contacts/complex.rs:178-405) and, optionally, the frame assembly of
//! `get_contacts` (src/contacts/mod.rs:61-214).  Layouts mirror the C structs field for field (`#[repr(C)]`).
//! How the reference's host code fills them: INTEGRATION.md.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

pub const ARP_OK: c_int = 0;
pub const ARP_ERR_IO: c_int = 1;
pub const ARP_ERR_BAD_GROUPS: c_int = 2; // utils.rs:77-80
pub const ARP_ERR_EMPTY_GROUP: c_int = 3; // utils.rs:109-112
pub const ARP_ERR_NO_RINGS: c_int = 4; // complex.rs:50
pub const ARP_ERR_ELEMENT: c_int = 5; // vdw.rs:24-28
pub const ARP_ERR_NARROWING: c_int = 6; // mod.rs:142-178
pub const ARP_ERR_RESIDUE_NAME: c_int = 7; // residues.rs:132
pub const ARP_ERR_ARG: c_int = 8;
pub const ARP_ERR_CUDA: c_int = 9;
pub const ARP_ERR_OOM: c_int = 10;
pub const ARP_ERR_LIMIT: c_int = 11;
pub const ARP_ERR_NCCL: c_int = 12;

pub const ARP_ATTR_ELEM_MASK: u32 = 0x0000_000F;
pub const ARP_ATTR_LIGAND: u32 = 0x0000_0010;
pub const ARP_ATTR_RECEPTOR: u32 = 0x0000_0020;
pub const ARP_ATTR_POS_CONF: u32 = 0x0000_0040;
pub const ARP_ATTR_NEG_CONF: u32 = 0x0000_0080;
pub const ARP_ATTR_POS_RES: u32 = 0x0000_0100;
pub const ARP_ATTR_DONOR: u32 = 0x0000_0200;
pub const ARP_ATTR_ACCEPTOR: u32 = 0x0000_0400;
pub const ARP_ATTR_WEAK_DONOR: u32 = 0x0000_0800;
pub const ARP_ATTR_HYDROPHOBIC: u32 = 0x0000_1000;
pub const ARP_ATTR_CYS_SG: u32 = 0x0000_2000;
pub const ARP_ATTR_GHOST: u32 = 0x0000_4000;
pub const ARP_RING_FLAG: u32 = 0x8000_0000;
pub const ARP_ENT_NARROW: u32 = 0x8000_0000;
pub const ARP_FLAG_EMIT_RADIUS_SCREEN: u32 = 0x1;
pub const ARP_FLAG_COUNT_CLASSIFIED: u32 = 0x2;
/// `Interaction` (structs.rs:6-51) in declaration order: `label as usize` indexes the enum.
pub const ARP_N_LABELS: usize = 19;
pub const ARP_STATIC_MAX_ATOMS: i64 = 131072;
pub const ARP_STATIC_MAX_CELLS: i64 = 262144;

#[repr(C)]
pub struct arp_structure {
    pub n_atoms: i64,
    pub n_residues: i64,
    pub n_rings: i64,
    pub n_h: i64,
    pub x: *const f64,
    pub y: *const f64,
    pub z: *const f64,
    pub attr: *const u32,
    pub chain_key: *const u32,
    pub res_ordinal: *const u32,
    pub res_id: *const u32,
    pub res_h_begin: *const u32,
    pub h_x: *const f64,
    pub h_y: *const f64,
    pub h_z: *const f64,
    pub res_cb: *const i32,
    pub res_sg: *const i32,
    pub ring_center: *const f64,
    pub ring_normal: *const f64,
    pub ring_attr: *const u32,
    pub ring_chain_key: *const u32,
    pub ring_res_ordinal: *const u32,
    pub bbox_min: [f64; 3],
    pub bbox_max: [f64; 3],
    pub has_bbox: i32,
    pub reserved_: i32,
    pub res_chain_key: *const u32,
    pub res_res_ordinal: *const u32,
    pub res_atom_begin: *const u32,
    pub x_milli: *const i32,
    pub y_milli: *const i32,
    pub z_milli: *const i32,
    pub atom_gid: *const u32,
    pub ring_gid: *const u32,
    pub res_tag: *const u32,
    pub xyz_milli21: *const u64,
    pub xyz_milli21_origin: [i32; 3],
    pub reserved2_: i32,
    pub attr16: *const u16,
}

#[repr(C)]
pub struct arp_params {
    pub vdw_comp: f64,
    pub dist_cutoff: f64,
    pub flags: u32,
    pub reserved_: u32,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct arp_row {
    pub from: u32,
    pub to: u32,
    pub label: u32,
    pub distance: f32,
}

#[repr(C)]
pub struct arp_result {
    pub n_rows: i64,
    pub rows: *const arp_row,
    pub n_structures: i64,
    pub atom_begin: *const i64,
    pub ring_begin: *const i64,
    pub pairs_screened: i64,
    pub pairs_in_cutoff: i64,
    pub pairs_classified: i64,
}

#[repr(C)]
pub struct arp_stats {
    pub n_atoms: i64,
    pub n_rings: i64,
    pub n_cells: i64,
    pub n_nonempty_cells: i64,
    pub n_rows: i64,
    pub kernel_launches: i64,
    pub pair_kernel_retries: i32,
    pub upload_copies: i32,
    pub n_screened_pairs: i64,
    pub n_filtered_pairs: i64,
    pub ms_h2d: f32,
    pub ms_grid: f32,
    pub ms_pairs: f32,
    pub ms_rings: f32,
    pub ms_d2h: f32,
    pub ms_run: f32,
    pub ms_screen: f32,
    pub ms_filter: f32,
    pub ms_exact: f32,
    pub reserved2_: f32,
    pub ms_wall_upload: f32,
    pub ms_wall_run: f32,
    pub ms_wall_download: f32,
    pub ms_wall_call: f32,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct arp_frame_side {
    pub chain: i32,
    pub resn: i32,
    pub resi: i32,
    pub icode: i32,
    pub altloc: i32,
    pub model: u32,
    pub chain_rank: u32,
    pub has_sc: u32,
}

#[repr(C)]
pub struct arp_frame_tables {
    pub n_atoms: i64,
    pub n_rings: i64,
    pub n_sides: i64,
    pub sides: *const arp_frame_side,
    pub side_sc: *const f64,
    pub ent_side: *const u32,
    pub ent_atomn: *const i32,
    pub ent_atomi: *const i32,
    pub ent_rank: *const u32,
    pub ent_tie: *const u32,
    pub label_rank: [u8; 32],
    pub n_chain_ranks: u32,
    pub n_ent_ranks: u32,
    pub max_tie: u32,
}

#[repr(C)]
pub struct arp_frame {
    pub n_rows: i64,
    pub model: *mut u32,
    pub interaction: *mut u8,
    pub distance: *mut f32,
    pub from_chain: *mut i32,
    pub from_resn: *mut i32,
    pub from_resi: *mut i32,
    pub from_insertion: *mut i32,
    pub from_altloc: *mut i32,
    pub from_atomn: *mut i32,
    pub from_atomi: *mut i32,
    pub to_chain: *mut i32,
    pub to_resn: *mut i32,
    pub to_resi: *mut i32,
    pub to_insertion: *mut i32,
    pub to_altloc: *mut i32,
    pub to_atomn: *mut i32,
    pub to_atomi: *mut i32,
    pub sc_centroid_dist: *mut f32,
    pub sc_dihedral: *mut f32,
    pub sc_centroid_angle: *mut f32,
    pub sc_valid: *mut u8,
    pub from: *mut u32,
    pub to: *mut u32,
    pub pairs_in_cutoff: i64,
    pub pairs_classified: i64,
    pub ms_sort: f32,
    pub ms_columns: f32,
    pub ms_d2h: f32,
    pub sort_passes: i32,
    pub key_bits: i32,
}

/// An `ncclUniqueId`: made by one rank (`arp_comm_get_id`), carried to the others by the host's own means.
#[repr(C)]
#[derive(Clone, Copy)]
pub struct arp_comm_id {
    pub bytes: [c_char; 128],
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct arp_slab_info {
    pub own_atoms: i64,
    pub own_residues: i64,
    pub own_rings: i64,
    pub ghost_atoms: i64,
    pub ghost_residues: i64,
    pub ghost_rings: i64,
    pub ghost_h: i64,
    pub sent_atoms: i64,
    pub sent_residues: i64,
    pub sent_rings: i64,
    pub bytes_sent: i64,
    pub bytes_received: i64,
    pub n_peers: i32,
    pub transport: i32,
    pub ms_select: f32,
    pub ms_transfer: f32,
    pub ms_unpack: f32,
    pub ms_total: f32,
}

/// Opaque context: one per (device, host thread).
#[repr(C)]
pub struct arp_ctx {
    _private: [u8; 0],
}

extern "C" {
    pub fn arp_label_name(label: c_int) -> *const c_char;
    pub fn arp_element_class(symbol: *const c_char) -> c_int;
    pub fn arp_element_radii(element_class: c_int, covalent_single: *mut f64, van_der_waals: *mut f64) -> c_int;
    pub fn arp_ctx_create(out: *mut *mut arp_ctx, device_ordinal: c_int) -> c_int;
    pub fn arp_ctx_destroy(ctx: *mut arp_ctx);
    pub fn arp_last_error(ctx: *const arp_ctx) -> *const c_char;
    pub fn arp_ctx_stream(ctx: *mut arp_ctx) -> *mut c_void;
    pub fn arp_contacts(ctx: *mut arp_ctx, s: *const arp_structure, p: *const arp_params, out: *mut *mut arp_result) -> c_int;
    pub fn arp_contacts_batch(ctx: *mut arp_ctx, s: *const *const arp_structure, n: i64, p: *const arp_params, out: *mut *mut arp_result) -> c_int;
    pub fn arp_result_free(ctx: *mut arp_ctx, r: *mut arp_result);
    pub fn arp_stage_upload(ctx: *mut arp_ctx, s: *const *const arp_structure, n: i64) -> c_int;
    pub fn arp_stage_run(ctx: *mut arp_ctx, p: *const arp_params) -> c_int;
    pub fn arp_stage_download(ctx: *mut arp_ctx, out: *mut *mut arp_result) -> c_int;
    pub fn arp_stage_stats(ctx: *mut arp_ctx, out: *mut arp_stats) -> c_int;
    pub fn arp_stage_neighbor_sum(ctx: *mut arp_ctx, weight: *const f32, select: *const u8, radius: f32, out: *mut f32) -> c_int;
    pub fn arp_neighbor_sum_stats(ctx: *mut arp_ctx, ms_grid: *mut f32, ms_sum: *mut f32, n_cells: *mut i64) -> c_int;
    pub fn arp_stage_frame(ctx: *mut arp_ctx, t: *const arp_frame_tables, out: *mut *mut arp_frame) -> c_int;
    pub fn arp_contacts_frame(ctx: *mut arp_ctx, s: *const arp_structure, p: *const arp_params, t: *const arp_frame_tables, out: *mut *mut arp_frame) -> c_int;
    pub fn arp_stage_frame_compact(ctx: *mut arp_ctx, t: *const arp_frame_tables, out: *mut *mut arp_frame) -> c_int;
    pub fn arp_contacts_frame_compact(ctx: *mut arp_ctx, s: *const arp_structure, p: *const arp_params, t: *const arp_frame_tables, out: *mut *mut arp_frame) -> c_int;
    pub fn arp_frame_free(ctx: *mut arp_ctx, f: *mut arp_frame);
    pub fn arp_comm_get_id(out: *mut arp_comm_id) -> c_int;
    pub fn arp_ctx_comm_init(ctx: *mut arp_ctx, id: *const arp_comm_id, rank: c_int, world: c_int) -> c_int;
    pub fn arp_ctx_create_multi(out: *mut *mut arp_ctx, devices: *const c_int, n_devices: c_int) -> c_int;
    pub fn arp_slab_exchange(ctx: *mut arp_ctx, p: *const arp_params, info: *mut arp_slab_info) -> c_int;
    pub fn arp_slab_exchange_multi(ctx: *const *mut arp_ctx, n: c_int, p: *const arp_params, info: *mut arp_slab_info) -> c_int;
    pub fn arp_neighbor_sum_static(device: c_int, stream: *mut c_void, x: *const f64, y: *const f64, z: *const f64, weight: *const f32, select: *const u8, n_atoms: i64, radius: f32, out: *mut f32) -> c_int;
    pub fn arp_host_alloc(out: *mut *mut c_void, bytes: i64) -> c_int;
    pub fn arp_host_free(p: *mut c_void);
}

/// The shape of the safe wrapper on the reference's side (complex.rs:178-186): rows are split by the kind of their ends.
/// `entry` turns (from, to, label, distance) back into a `ResultEntry` through the index -> hierarchy vector the host
/// kept while it filled `arp_structure`.
pub fn row_kind(r: &arp_row) -> (bool, bool) {
    (r.from & ARP_RING_FLAG != 0, r.to & ARP_RING_FLAG != 0)
}
