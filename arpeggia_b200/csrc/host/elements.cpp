// Element classes / radii and label names shared by host and device code.
#include <cstring>

#include "common.hpp"

namespace arp {
std::string& last_error() {
  thread_local std::string e;
  return e;
}
}  // namespace arp

namespace {
// pdbtbx Element::atomic_radius(): covalent_single (Pyykko 2009) and van_der_waals (Alvarez 2013), angstrom
// (SURVEY.md 8.2).  Class index = position in this table; class 0 must stay hydrogen.
struct ElemRow {
  const char* sym;
  double cov, vdw;
};
const ElemRow kElems[] = {
    {"H", 0.32, 1.20},  {"C", 0.75, 1.77},  {"N", 0.71, 1.66},  {"O", 0.63, 1.50},  {"S", 1.03, 1.89},  {"P", 1.11, 1.90},
    {"SE", 1.16, 1.82}, {"F", 0.64, 1.46},  {"CL", 0.99, 1.82}, {"BR", 1.14, 1.86}, {"I", 1.33, 2.04},
};
const int kNumElems = (int)(sizeof kElems / sizeof kElems[0]);

const char* kLabelNames[ARP_N_LABELS] = {
    "StericClash", "CovalentBond", "Disulfide", "VanDerWaalsContact", "IonicBond", "HydrogenBond", "WeakHydrogenBond",
    "PolarContact", "WeakPolarContact", "IonicRepulsion", "SaltBridge", "PiDisplacedStacking", "PiTStacking",
    "PiSandwichStacking", "PiParallelInPlaneStacking", "PiTiltedStacking", "PiLStacking", "CationPi", "HydrophobicContact"};
}  // namespace

extern "C" {

const char* arp_label_name(int label) { return (label >= 0 && label < ARP_N_LABELS) ? kLabelNames[label] : ""; }

int arp_element_class(const char* symbol) {
  if (!symbol) return -1;
  for (int k = 0; k < kNumElems; ++k)
    if (std::strcmp(symbol, kElems[k].sym) == 0) return k;
  return -1;
}

int arp_element_radii(int cls, double* cov, double* vdw) {
  if (cls < 0 || cls >= kNumElems) return ARP_ERR_ARG;
  if (cov) *cov = kElems[cls].cov;
  if (vdw) *vdw = kElems[cls].vdw;
  return ARP_OK;
}

const char* arph_last_error(void) { return arp::last_error().c_str(); }
}
