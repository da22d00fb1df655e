// arph_contacts / arph_contacts_file: the host mirror of get_contacts (src/contacts/mod.rs:61-136) and of
// arpeggia.contacts() (src/python.rs:31-56): load -> filters -> pack -> arp_contacts (device) -> table.
#include <chrono>

#include "packed.hpp"

namespace {
double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}  // namespace

extern "C" {

int arph_contacts(arp_ctx* ctx, const arp_atom_record* atoms, int64_t n, const char* groups, double vdw_comp, double dist_cutoff,
                  int ignore_zero_occupancy, uint32_t flags, arph_table** out) {
  *out = nullptr;
  if (!ctx) {
    arp::last_error() = "arph_contacts needs a device context (there is no CPU fallback)";
    return ARP_ERR_CUDA;
  }
  const double t0 = now_s();
  arph_packed* pk = nullptr;
  int rc = arph_pack(atoms, n, groups, ignore_zero_occupancy, flags, &pk);
  if (rc != ARP_OK) return rc;
  const double t1 = now_s();
  arp_params prm{};
  prm.vdw_comp = vdw_comp;
  prm.dist_cutoff = dist_cutoff;
  // the table reports both workload counters, unless the caller only wants the frame the reference returns
  prm.flags = (flags & ARPH_FLAG_NO_COUNTERS) ? ARP_FLAG_EMIT_RADIUS_SCREEN : ARP_FLAG_COUNT_CLASSIFIED;
  // rows stay on the device: sorted and joined there (arp_stage_frame_compact); 26 bytes per row come back and the
  // per-entity columns are gathered from the side tables on the host's threads (ARPH_FLAG_DEVICE_COLUMNS: on the device)
  arp::FrameTables ft;
  arp_frame* frame = nullptr;
  double t2 = t1, t3 = t1;
  try {
    arp::build_frame_tables(arph_packed_ref(pk), ft);
    t2 = now_s();
    rc = (flags & ARPH_FLAG_DEVICE_COLUMNS) ? arp_contacts_frame(ctx, arph_packed_structure(pk), &prm, &ft.t, &frame)
                                            : arp_contacts_frame_compact(ctx, arph_packed_structure(pk), &prm, &ft.t, &frame);
    if (rc != ARP_OK) {
      arp::last_error() = arp_last_error(ctx);
      arph_packed_free(pk);
      return rc;
    }
    t3 = now_s();
    *out = arp::table_from_frame(arph_packed_ref(pk), frame, &ft.t);
    frame = nullptr;
    (*out)->t_setup_s = t2 - t0;
    (*out)->t_classify_s = t3 - t2;
    (*out)->t_table_s = now_s() - t3;
  } catch (const arp::Error& e) {
    arp::last_error() = e.msg;
    rc = e.code;
  } catch (const std::exception& e) {
    arp::last_error() = e.what();
    rc = ARP_ERR_OOM;
  }
  if (frame) arp_frame_free(ctx, frame);
  arph_packed_free(pk);
  return rc;
}

int arph_contacts_file(arp_ctx* ctx, const char* path, const char* groups, double vdw_comp, double dist_cutoff,
                       int ignore_zero_occupancy, uint32_t flags, arph_table** out) {
  *out = nullptr;
  arp_atom_record* recs = nullptr;
  int64_t n = 0;
  int rc = arph_parse_file(path, &recs, &n);
  if (rc != ARP_OK) return rc;
  rc = arph_contacts(ctx, recs, n, groups, vdw_comp, dist_cutoff, ignore_zero_occupancy, flags, out);
  arph_free(recs);
  return rc;
}
}
