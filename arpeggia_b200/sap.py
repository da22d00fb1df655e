"""The neighbour half of the SAP score (src/sap.rs:181-215) on the device: SURVEY.md 8(f) rank 3.

    SAP(i) = sum over side-chain atoms j within R of side-chain atom i (i included) of
             hydrophobicity(resn_j) * clamp(sasa_j / max_sc_asa(resn_j), 0, 1)

The per-atom SASA comes from the caller (the reference computes it with the rust-sasa crate, src/sasa.rs: a different
analysis, outside this repository's path); the two residue tables and the radius query are restated here, the query runs
through `arp_stage_neighbor_sum` (cell list + exact FP64 distance test on the GPU, no CPU fallback).

Unpinned (pdbtbx 0.12 source is not under /root/reference): `is_sidechain()` / `is_backbone()`.  pdbtbx documents the
backbone as the atoms named N, CA, C, O of an amino acid (plus their hydrogens, which sap.rs:177 strips anyway);
`sidechain_mask` follows that.  The structure is staged through the contacts packer, which needs at least one aromatic
residue (complex.rs:50 panics otherwise); `Context.neighbor_sum` itself has no such requirement.
"""
from __future__ import annotations

import numpy as np

from .core import Context, Packed

_F = np.float32
_GLY = _F(0.501)
# Black & Mould (1991) minus glycine, f32 like the reference's literals (sap.rs:42-67)
HYDROPHOBICITY = {
    "ALA": _F(0.616) - _GLY, "ARG": _F(0.000) - _GLY, "ASN": _F(0.236) - _GLY, "ASP": _F(0.028) - _GLY, "CYS": _F(0.680) - _GLY,
    "GLU": _F(0.043) - _GLY, "GLN": _F(0.251) - _GLY, "GLY": _F(0.000), "HIS": _F(0.165) - _GLY, "ILE": _F(0.943) - _GLY,
    "LEU": _F(0.943) - _GLY, "LYS": _F(0.283) - _GLY, "MET": _F(0.738) - _GLY, "PHE": _F(1.000) - _GLY, "PRO": _F(0.711) - _GLY,
    "SER": _F(0.359) - _GLY, "THR": _F(0.450) - _GLY, "TRP": _F(0.878) - _GLY, "TYR": _F(0.880) - _GLY, "VAL": _F(0.825) - _GLY,
}
# maximum side-chain SASA in A^2 (Rosetta sap_sasa_calib.dat; sap.rs:77-101)
SC_MAX_ASA = {
    "ALA": 15.395, "ARG": 124.338, "ASN": 90.303, "ASP": 87.601, "CYS": 46.456, "GLN": 99.186, "GLY": 3.229, "GLU": 95.534,
    "HIS": 96.532, "ILE": 31.448, "LEU": 30.271, "LYS": 61.962, "MET": 65.233, "PHE": 67.945, "PRO": 17.812, "SER": 39.355,
    "THR": 42.648, "TRP": 101.491, "TYR": 94.478, "VAL": 26.702,
}
SOLVENT_RESIDUES = ("HOH", "H2O", "D2O", "WAT", "TIP", "TIP3", "TIP4", "SPC")  # sasa.rs:45
ION_RESIDUES = ("NA", "CL", "K", "CA", "MG", "ZN", "FE", "MN", "CU", "CO", "NI", "CD", "SO4", "PO4", "NO3", "ACE", "NH2")  # sasa.rs:48-51
BACKBONE_NAMES = ("N", "CA", "C", "O")


def get_hydrophobicity(resn: str):
    """sap.rs:42-67: None for anything but the 20 amino acids (case-insensitive)."""
    return HYDROPHOBICITY.get(resn.upper())


def get_sc_max_asa(resn: str):
    """sap.rs:77-101."""
    v = SC_MAX_ASA.get(resn.upper())
    return None if v is None else _F(v)


def prepare_atoms(atoms: np.ndarray, chains: str = "") -> np.ndarray:
    """prepare_pdb_for_sasa(pdb, true, true, chains), sasa.rs:112-140: chain filter, hydrogens out, solvent / ion
    residues out.  Returns the kept records (file order)."""
    keep = np.ones(len(atoms), dtype=bool)
    ids = {c.strip() for c in chains.split(",") if c.strip()} if chains else set()
    if ids:
        keep &= np.isin(atoms["chain"], [c.encode() for c in ids])
    keep &= atoms["element"] != b"H"
    keep &= ~np.isin(atoms["resn"], [r.encode() for r in SOLVENT_RESIDUES + ION_RESIDUES])
    return atoms[keep]


def sidechain_mask(atoms: np.ndarray) -> np.ndarray:
    """h.is_sidechain() of pdbtbx (unpinned, see the module docstring): an amino-acid atom that is not N / CA / C / O."""
    amino = np.isin(atoms["resn"], [r.encode() for r in SC_MAX_ASA])
    return amino & ~np.isin(atoms["name"], [b.encode() for b in BACKBONE_NAMES])


def sap_weights(atoms: np.ndarray, sasa: np.ndarray) -> np.ndarray:
    """The contribution of every atom as a neighbour, sap.rs:199-210, in f32 like the reference:
    hydrophobicity(resn) * clamp(sasa / max_sc_asa(resn), 0, 1); 0 where the residue has no table entry."""
    sasa = np.asarray(sasa, dtype=np.float32)
    w = np.zeros(len(atoms), dtype=np.float32)
    for resn, h in HYDROPHOBICITY.items():
        m = np.char.upper(atoms["resn"]) == resn.encode()
        if m.any():
            w[m] = h * np.clip(sasa[m] / _F(SC_MAX_ASA[resn]), _F(0), _F(1))
    return w


def per_atom_sap(atoms: np.ndarray, sasa: np.ndarray, sap_radius: float = 5.0, *, ctx: Context | None = None, device: int = 0) -> dict:
    """get_per_atom_sap_score (sap.rs:124-262) given the per-atom SASA of the prepared atoms: returns the side-chain atoms'
    serial numbers, their SASA and SAP score, sorted by serial (sap.rs:259).  `atoms` are the records that
    prepare_atoms() returned (one conformer per atom: altloc handling belongs to the SASA step)."""
    from .core import default_context

    ctx = ctx or default_context(device)
    sel = sidechain_mask(atoms)
    w = sap_weights(atoms, sasa)
    packed = Packed(atoms, "/", flags=2)  # records are already filtered; rings / groups are irrelevant here
    n = packed.structure.n_atoms
    if n != len(atoms):
        raise ValueError("per_atom_sap needs one record per atom (no alternate conformers): resolve altlocs before the SASA step")
    # the packer walks the hierarchy (chain, residue, conformer, atom): map the records onto its order by serial number
    rec_of = {int(sn): k for k, sn in enumerate(atoms["serial"])}
    if len(rec_of) != len(atoms):
        raise ValueError("per_atom_sap needs unique atom serial numbers")
    perm = np.array([rec_of[int(sn)] for sn in packed.atom_serials()], dtype=np.int64)  # packed atom -> record
    ctx.upload(packed)
    packed_score = ctx.neighbor_sum(w[perm], sel[perm], float(sap_radius))
    score = np.zeros(len(atoms), dtype=np.float32)
    score[perm] = packed_score
    order = np.argsort(atoms["serial"][sel], kind="stable")
    return {"atomi": atoms["serial"][sel][order], "sasa": np.asarray(sasa, dtype=np.float32)[sel][order], "sap_score": score[sel][order]}
