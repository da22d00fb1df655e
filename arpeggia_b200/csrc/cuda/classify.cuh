// Per-pair filter and interaction classification, in registers, exact FP64 (each op rounded once, no FMA).
// Restates (file:line relative to the reference):
//   should_compare_entities / should_compare_residues   src/contacts/complex.rs:76-131
//   classification closure of get_atomic_contacts        src/contacts/complex.rs:217-296
//   find_vdw_contact / is_disulfide                      src/contacts/vdw.rs:19-80
//   find_ionic_bond / find_ionic_repulsion               src/contacts/ionic.rs:11-35
//   find_hydrogen_bond / find_weak_hydrogen_bond         src/contacts/hbond.rs:30-110
//   find_hydrophobic_contact                             src/contacts/hydrophobic.rs:10-24
//   find_cation_pi / find_pi_pi, Plane angle helpers     src/contacts/aromatic.rs:14-64, residues.rs:31-75
// Angle comparisons run in the cosine domain against host-computed boundary doubles (DevAngles).
#pragma once
#include "dev_common.cuh"

namespace arpd {

#define LBIT(l) (1u << (l))

// complex.rs:76-131 for the ordered pair (p -> q), symmetric = true, plus the outer-loop conditions
// complex.rs:200-202 (p in ligand, p not H) and :205 (q in receptor).  Ghost atoms never start a row.
__device__ __forceinline__ bool pair_filter(uint32_t wp, uint32_t kp, uint32_t op, uint32_t wq, uint32_t kq, uint32_t oq) {
  if (!(wp & ARP_ATTR_LIGAND) || !(wq & ARP_ATTR_RECEPTOR)) return false;
  if ((wp & ARP_ATTR_ELEM_MASK) == 0u || (wq & ARP_ATTR_ELEM_MASK) == 0u) return false;  // hydrogens, complex.rs:83-87
  if (wp & ARP_ATTR_GHOST) return false;
  if ((kp ^ kq) >> 24) return false;                                                    // different models, complex.rs:96
  if (kp == kq) return (oq > 1u) && (op < oq - 1u);                                       // complex.rs:112-113
  const bool both = (wp & wq & ARP_ATTR_LIGAND) && (wp & wq & ARP_ATTR_RECEPTOR);
  return !(both && kp > kq);                                                              // complex.rs:124-129
}

// Both orientations of an unordered pair at once: 0 = neither passes, 1 = a -> b, 2 = b -> a.  At most one orientation
// can pass: within a chain the row runs from the earlier residue to the later one (complex.rs:112-113); across chains
// either only one side satisfies ligand -> receptor, or both do and the string-order dedupe keeps one (complex.rs:124-129).
__device__ __forceinline__ int pair_direction(uint32_t wa, uint32_t ka, uint32_t oa, uint32_t wb, uint32_t kb, uint32_t ob) {
  if ((wa & ARP_ATTR_ELEM_MASK) == 0u || (wb & ARP_ATTR_ELEM_MASK) == 0u) return 0;  // hydrogens, complex.rs:83-87
  if ((ka ^ kb) >> 24) return 0;                                                      // different models, complex.rs:96
  const bool ab = (wa & ARP_ATTR_LIGAND) && (wb & ARP_ATTR_RECEPTOR) && !(wa & ARP_ATTR_GHOST);
  const bool ba = (wb & ARP_ATTR_LIGAND) && (wa & ARP_ATTR_RECEPTOR) && !(wb & ARP_ATTR_GHOST);
  if (ka == kb) {
    if (ab && ob > 1u && oa < ob - 1u) return 1;
    if (ba && oa > 1u && ob < oa - 1u) return 2;
    return 0;
  }
  const bool both = (wa & wb & ARP_ATTR_LIGAND) && (wa & wb & ARP_ATTR_RECEPTOR);
  if (ab && !(both && ka > kb)) return 1;
  if (ba && !(both && kb > ka)) return 2;
  return 0;
}

// complex.rs:94-131 with symmetric = false (ring -> atom), ring r, atom y
__device__ __forceinline__ bool ring_atom_filter(uint32_t wr, uint32_t kr, uint32_t orr, uint32_t wy, uint32_t ky, uint32_t oy) {
  if (wr & ARP_ATTR_GHOST) return false;
  if ((kr ^ ky) >> 24) return false;
  const bool lr = wr & ARP_ATTR_LIGAND, rr = wr & ARP_ATTR_RECEPTOR, ly = wy & ARP_ATTR_LIGAND, ry = wy & ARP_ATTR_RECEPTOR;
  if (!((lr && ry) || (ly && rr))) return false;
  if (kr == ky) {
    const bool near = (oy == orr) || (oy == orr + 1u) || (orr > 0u && oy == orr - 1u);  // complex.rs:115-119
    return !near;
  }
  return true;
}

// vdw.rs:46-80.  ra / rb = residue ids of the two SG atoms.
__device__ __forceinline__ bool is_disulfide(const PairArgs& G, const DevParams& P, uint32_t ra, uint32_t rb) {
  const int cb1 = G.res_cb[ra], s1 = G.res_sg[ra], cb2 = G.res_cb[rb], s2 = G.res_sg[rb];
  if (cb1 < 0 || cb2 < 0 || s1 < 0 || s2 < 0) {  // `.find(..).unwrap()` panics in the reference
    atomicOr(&G.ctr->error, 1u);
    return false;
  }
  // pdbtbx Atom::dihedral(cb1; s1, s2, cb2): angle between the normals of (cb1,s1,s2) and (s1,s2,cb2)
  const double ax = G.x[cb1], ay = G.y[cb1], az = G.z[cb1];
  const double bx = G.x[s1], by = G.y[s1], bz = G.z[s1];
  const double cx = G.x[s2], cy = G.y[s2], cz = G.z[s2];
  const double dx = G.x[cb2], dy = G.y[cb2], dz = G.z[cb2];
  const double bax = dsub(ax, bx), bay = dsub(ay, by), baz = dsub(az, bz);
  const double bcx = dsub(cx, bx), bcy = dsub(cy, by), bcz = dsub(cz, bz);
  const double cbx = dsub(bx, cx), cby = dsub(by, cy), cbz = dsub(bz, cz);
  const double cdx = dsub(dx, cx), cdy = dsub(dy, cy), cdz = dsub(dz, cz);
  const double n1x = dsub(dmul(bay, bcz), dmul(baz, bcy)), n1y = dsub(dmul(baz, bcx), dmul(bax, bcz)),
               n1z = dsub(dmul(bax, bcy), dmul(bay, bcx));
  const double n2x = dsub(dmul(cby, cdz), dmul(cbz, cdy)), n2y = dsub(dmul(cbz, cdx), dmul(cbx, cdz)),
               n2z = dsub(dmul(cbx, cdy), dmul(cby, cdx));
  const double c = __ddiv_rn(dot3(n1x, n1y, n1z, n2x, n2y, n2z), dmul(norm3(n1x, n1y, n1z), norm3(n2x, n2y, n2z)));
  // (60.0..=120.0).contains(&deg(acos(c)).abs())
  return c <= P.ang.a_ge_60 && c >= P.ang.a_le_120;
}

// shared body of hbond.rs:30-66 / :74-110.  D = donor position, A = acceptor position, `res` = donor residue.
// Returns 2 = (weak) hydrogen bond, 1 = (weak) polar contact, 0 = nothing.
__device__ __forceinline__ int hbond_body(const PairArgs& G, const DevParams& P, double Dx, double Dy, double Dz, double Ax, double Ay,
                                          double Az, uint32_t res, uint32_t acc_class, double dist, double cos_bound) {
  if (dist <= 4.0) {
    const double reach = dadd(dadd(P.vdw[0], P.vdw[acc_class]), P.comp);  // h_vdw + acceptor_vdw + vdw_comp_factor
    const uint32_t hb = G.res_h_begin[res], he = G.res_h_begin[res + 1];
    for (uint32_t h = hb; h < he; ++h) {
      const double Hx = G.hx[h], Hy = G.hy[h], Hz = G.hz[h];
      const double ex = dsub(Ax, Hx), ey = dsub(Ay, Hy), ez = dsub(Az, Hz);  // h.distance(acceptor)
      if (!(norm3(ex, ey, ez) <= reach)) continue;
      // donor.angle(h, acceptor): angle at h between h->donor and h->acceptor
      const double fx = dsub(Dx, Hx), fy = dsub(Dy, Hy), fz = dsub(Dz, Hz);
      const double c = __ddiv_rn(dot3(fx, fy, fz, ex, ey, ez), dmul(norm3(fx, fy, fz), norm3(ex, ey, ez)));
      if (c <= cos_bound && c >= -1.0) return 2;
    }
  }
  return dist <= 3.5 ? 1 : 0;
}

// Labels of the ordered pair (e1 = a, e2 = b) as a bit mask over arp_label.  complex.rs:217-296.
// sa / sb: residue ids (arp_structure.res_id) of the two atoms
__device__ __forceinline__ uint32_t classify_pair(const PairArgs& G, const DevParams& P, const RecB& a, uint32_t wa, uint32_t sa,
                                                  const RecB& b, uint32_t wb, uint32_t sb, double dist) {
  const uint32_t ea = wa & ARP_ATTR_ELEM_MASK, eb = wb & ARP_ATTR_ELEM_MASK;
  const double sum_cov = dadd(P.cov[ea], P.cov[eb]);
  const double sum_vdw = dadd(P.vdw[ea], P.vdw[eb]);
  uint32_t m = 0;
  if (dist < dsub(sum_cov, P.comp)) return LBIT(ARP_StericClash);  // complex.rs:233-235: a clash suppresses the rest
  if (dist < dadd(sum_cov, P.comp)) {
    const bool ss = (wa & wb & ARP_ATTR_CYS_SG) && is_disulfide(G, P, sa, sb);
    m |= ss ? LBIT(ARP_Disulfide) : LBIT(ARP_CovalentBond);
  } else if (dist < dadd(sum_vdw, P.comp)) {
    m |= LBIT(ARP_VanDerWaalsContact);
  }
  // ionic.rs:11-22
  const bool ion_pair = ((wa & ARP_ATTR_POS_CONF) && (wb & ARP_ATTR_NEG_CONF)) || ((wb & ARP_ATTR_POS_CONF) && (wa & ARP_ATTR_NEG_CONF));
  const bool ionic = ion_pair && dist <= 4.0;
  // hbond.rs:30-66 (k = 0, donors) and hbond.rs:74-110 (k = 1, weak donors), both with is_donor_acceptor_pair: entity1 is
  // tried as the donor first.  One rolled loop = one instance of hbond_body (instruction-cache footprint).
  int hb = 0, wk = 0;
#pragma unroll 1
  for (int k = 0; k < 2; ++k) {
    const uint32_t dflag = k ? ARP_ATTR_WEAK_DONOR : ARP_ATTR_DONOR;
    const bool fa = (wa & dflag) && (wb & ARP_ATTR_ACCEPTOR);
    const bool fb = !fa && (wb & dflag) && (wa & ARP_ATTR_ACCEPTOR);
    if (!(fa || fb)) continue;
    const int r = hbond_body(G, P, fa ? a.x : b.x, fa ? a.y : b.y, fa ? a.z : b.z, fa ? b.x : a.x, fa ? b.y : a.y, fa ? b.z : a.z,
                             fa ? sa : sb, fa ? eb : ea, dist, k ? P.ang.a_ge_130 : P.ang.a_ge_90);
    if (k) wk = r; else hb = r;
  }
  // complex.rs:240-251
  if (ionic)
    m |= (hb == 2) ? LBIT(ARP_SaltBridge) : LBIT(ARP_IonicBond);
  else if (hb == 2)
    m |= LBIT(ARP_HydrogenBond);
  else if (hb == 1)
    m |= LBIT(ARP_PolarContact);
  if (wk == 2) m |= LBIT(ARP_WeakHydrogenBond);
  if (wk == 1) m |= LBIT(ARP_WeakPolarContact);
  // ionic.rs:25-35
  const bool same = (wa & wb & ARP_ATTR_POS_CONF) || (wa & wb & ARP_ATTR_NEG_CONF);
  if (same && dist <= 4.0) m |= LBIT(ARP_IonicRepulsion);
  // hydrophobic.rs:10-24
  if ((wa & wb & ARP_ATTR_HYDROPHOBIC) && dist <= 4.5) m |= LBIT(ARP_HydrophobicContact);
  return m;
}

// aromatic.rs:33-64.  Returns the label or -1.  c1/n1 = ring1 plane, c2/n2 = ring2 plane; dist out.
__device__ __forceinline__ int find_pi_pi(const DevParams& P, const double* c1, const double* n1, const double* c2, const double* n2,
                                          double* dist_out) {
  const double ax = dsub(c1[0], c2[0]), ay = dsub(c1[1], c2[1]), az = dsub(c1[2], c2[2]);
  const double dist = norm3(ax, ay, az);
  *dist_out = dist;
  if (!(dist <= 6.0)) return -1;
  // theta = ring1.point_vec_angle(ring2.center): v = c2 - c1
  const double vx = dsub(c2[0], c1[0]), vy = dsub(c2[1], c1[1]), vz = dsub(c2[2], c1[2]);
  const double n1n = norm3(n1[0], n1[1], n1[2]);
  const double ct = __ddiv_rn(dot3(n1[0], n1[1], n1[2], vx, vy, vz), dmul(n1n, norm3(vx, vy, vz)));
  const double cd = __ddiv_rn(dot3(n1[0], n1[1], n1[2], n2[0], n2[1], n2[2]), dmul(n1n, norm3(n2[0], n2[1], n2[2])));
  const DevAngles& A = P.ang;
  const bool d_ok = cd >= -1.0 && cd <= 1.0, t_ok = ct >= -1.0 && ct <= 1.0;  // otherwise acos -> NaN, every arm fails
  if (!d_ok) return -1;
  if (fold_pred(cd, A.f_le_30_pos, A.f_le_30_neg)) {
    if (fold_pred(ct, A.f_le_30_pos, A.f_le_30_neg)) return ARP_PiSandwichStacking;
    if (fold_pred(ct, A.f_le_60_pos, A.f_le_60_neg)) return ARP_PiDisplacedStacking;
    if (t_ok) return ARP_PiParallelInPlaneStacking;  // folded theta is always <= 90
    return -1;
  }
  if (fold_pred(cd, A.f_le_60_pos, A.f_le_60_neg)) return ARP_PiTiltedStacking;
  // dihedral <= 90 holds for every valid cosine
  const bool t_ge_30 = t_ok && !fold_pred(ct, A.f_lt_30_pos, A.f_lt_30_neg);
  const bool t_lt_60 = fold_pred(ct, A.f_lt_60_pos, A.f_lt_60_neg);
  if (t_ge_30 && t_lt_60) return ARP_PiLStacking;
  return dist <= 5.0 ? ARP_PiTStacking : -1;
}

}  // namespace arpd
