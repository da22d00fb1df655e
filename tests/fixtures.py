"""Synthetic edge-case structures for behaviour no reference fixture exercises ("unpinned"):
explicit hydrogens (H-bond / weak H-bond angle branches, SaltBridge), alternate locations, insertion codes,
several models, zero occupancy, steric clashes and exact cutoff ties.  Deterministic, no RNG."""
import numpy as np

from oracle.oracle import ATOM_DTYPE


def _unit(v):
    n = np.linalg.norm(v)
    return v / n if n > 0 else np.array([1.0, 0.0, 0.0])


def protonate_crude(atoms: np.ndarray, every: int = 1) -> np.ndarray:
    """Add one hydrogen to every N / O / S / C atom that has a bonded heavy neighbour in its residue: 1.0 A away,
    pointing away from the neighbours' centroid, snapped to the 0.001 A lattice.  Chemically crude, geometrically
    varied: gives D-H...A angles on both sides of the 90 / 130 degree thresholds."""
    out = []
    serial = int(atoms["serial"].max()) + 1
    keys = list(zip(atoms["model"].tolist(), atoms["chain"].tolist(), atoms["resi"].tolist(), atoms["icode"].tolist()))
    start = 0
    n = len(atoms)
    while start < n:
        end = start
        while end < n and keys[end] == keys[start]:
            end += 1
        res = atoms[start:end]
        xyz = np.stack([res["x"], res["y"], res["z"]], 1)
        out.append(res)
        if res["resn"][0] != b"HOH":
            hs = []
            for i in range(len(res)):
                if (start + i) % every:
                    continue
                d = np.linalg.norm(xyz - xyz[i], axis=1)
                nb = (d > 0.1) & (d < 1.9)
                if not nb.any():
                    continue
                pos = xyz[i] + 1.0 * _unit(xyz[i] - xyz[nb].mean(0))
                h = res[i:i + 1].copy()
                h["x"], h["y"], h["z"] = (np.rint(pos * 1000) / 1000.0)
                h["name"] = b"H" + res["name"][i][:3]
                h["element"] = b"H"
                h["serial"] = serial
                serial += 1
                hs.append(h)
            if hs:
                out.append(np.concatenate(hs))
        start = end
    return np.concatenate(out).astype(ATOM_DTYPE)


def with_altlocs(atoms: np.ndarray, resis=(6, 27, 45, 59)) -> np.ndarray:
    """Side chains of the given residues get altloc A and a copy B shifted by 0.3 A; backbone stays blank."""
    out = []
    serial = int(atoms["serial"].max()) + 1
    for r in atoms:
        r = r.copy()
        side = r["resi"] in resis and r["name"] not in (b"N", b"CA", b"C", b"O") and r["resn"] != b"HOH"
        if side:
            r["altloc"] = b"A"
            out.append(r)
            b = r.copy()
            b["altloc"] = b"B"
            b["x"] = round(float(b["x"]) + 0.3, 3)
            b["z"] = round(float(b["z"]) - 0.2, 3)
            b["serial"] = serial
            serial += 1
            out.append(b)
        else:
            out.append(r)
    return np.array(out, dtype=ATOM_DTYPE)


def with_insertion_codes(atoms: np.ndarray) -> np.ndarray:
    """Residues 30..33 become 30, 30A, 30B, 30C (same resi, different icode)."""
    out = atoms.copy()
    for k, code in zip((31, 32, 33), (b"A", b"B", b"C")):
        m = (out["resi"] == k) & (out["resn"] != b"HOH")
        out["resi"][m] = 30
        out["icode"][m] = code
    return out


def two_models(atoms: np.ndarray) -> np.ndarray:
    """Model 1 = the structure, model 2 = the structure rotated by 90 degrees about z and stretched 10 % along z
    (not a rigid motion, so the side-chain planes of the two models differ)."""
    a = atoms.copy()
    a["model"] = 1
    b = atoms.copy()
    b["model"] = 2
    x, y = b["x"].copy(), b["y"].copy()
    b["x"], b["y"] = np.rint(-y * 1000) / 1000.0, np.rint(x * 1000) / 1000.0
    b["z"] = np.rint(b["z"] * 1100) / 1000.0
    return np.concatenate([a, b])


def with_zero_occupancy(atoms: np.ndarray) -> np.ndarray:
    out = atoms.copy()
    out["occupancy"][(out["resi"] % 7 == 3) & (out["name"] != b"CA")] = 0.0
    return out


def cutoff_ties(atoms: np.ndarray) -> np.ndarray:
    """Extra waters placed so that squared distances hit rc^2 = 42.25 exactly (6.5,0,0), (3.9,5.2,0), (2.5,6,0)
    and sit one lattice step either side of it; plus a pair closer than the clash threshold."""
    base = atoms[atoms["name"] == b"CA"][0]
    o = np.array([200.0, 200.0, 200.0])
    offs = [(0, 0, 0), (6.5, 0, 0), (3.9, 5.2, 0), (2.5, 6.0, 0), (0, 6.501, 0), (0, 0, 6.499), (0, -6.5, 0), (0.9, 0, 0),
            (-6.5, 0, 0), (-3.9, -5.2, 0), (0, 0, -6.5)]
    extra = np.zeros(len(offs), dtype=ATOM_DTYPE)
    for k, d in enumerate(offs):
        e = base.copy()
        p = o + np.array(d, dtype=float)
        e["x"], e["y"], e["z"] = p
        e["resn"], e["name"], e["element"], e["hetero"] = b"HOH", b"O", b"O", 1
        e["chain"] = b"W"
        e["resi"] = 900 + 3 * k  # ordinals far enough apart for the same-chain rule
        e["serial"] = 90000 + k
        extra[k] = e
    # fillers so that consecutive tie waters are not ordinal neighbours
    fill = np.zeros(3 * len(offs), dtype=ATOM_DTYPE)
    for k in range(len(fill)):
        e = extra[0].copy()
        e["x"] = 400.0 + 20.0 * k
        e["resi"] = 900 + k
        e["serial"] = 91000 + k
        fill[k] = e
    keep = ~np.isin(fill["resi"], extra["resi"])
    allw = np.concatenate([extra, fill[keep]])
    allw = allw[np.argsort(allw["resi"], kind="stable")]
    return np.concatenate([atoms, allw])


def with_salt_bridges(atoms: np.ndarray) -> np.ndarray:
    """Chain S: LYS NZ-H...OD1 ASP pairs at 2.8 A with the hydrogen on the axis (SaltBridge, complex.rs:241-247),
    off the axis at 80 degrees (stays IonicBond), and a GLU/GLU like-charge pair (IonicRepulsion)."""
    base = atoms[atoms["name"] == b"CA"][0]
    recs = []

    def add(resn, resi, name, el, p):
        e = base.copy()
        # partners live in chain T so the same-chain neighbour rule (complex.rs:112-113) does not hide them
        e["chain"] = b"T" if resi % 10 == 5 else b"S"
        e["resn"], e["resi"], e["name"], e["element"] = resn, resi, name, el
        e["x"], e["y"], e["z"] = p
        e["serial"] = 80000 + len(recs)
        recs.append(e)

    for k, (hx, hy) in enumerate([(1.0, 0.0), (0.1736, 0.9848)]):  # D-H...A angle 180 and ~100 -> both >= 90; see below
        o = np.array([300.0 + 30.0 * k, 300.0, 300.0])
        add(b"LYS", 10 + 10 * k, b"NZ", b"N", o)
        add(b"LYS", 10 + 10 * k, b"HZ1", b"H", o + np.array([hx, hy, 0.0]))
        add(b"ASP", 15 + 10 * k, b"OD1", b"O", o + np.array([2.8, 0.0, 0.0]))
    o = np.array([300.0, 340.0, 300.0])  # hydrogen pointing away: angle < 90 -> ionic bond only
    add(b"LYS", 40, b"NZ", b"N", o)
    add(b"LYS", 40, b"HZ1", b"H", o + np.array([-1.0, 0.0, 0.0]))
    add(b"ASP", 45, b"OD2", b"O", o + np.array([2.8, 0.0, 0.0]))
    o = np.array([300.0, 380.0, 300.0])
    add(b"GLU", 50, b"OE1", b"O", o)
    add(b"GLU", 55, b"OE2", b"O", o + np.array([3.7, 0.0, 0.0]))
    return np.concatenate([atoms, np.array(recs, dtype=ATOM_DTYPE)])


def all_cases(ubq: np.ndarray) -> dict:
    return {
        "protonated": protonate_crude(ubq),
        "salt_bridges": with_salt_bridges(protonate_crude(ubq, every=3)),
        "altloc": with_altlocs(ubq),
        "altloc_protonated": with_altlocs(protonate_crude(ubq, every=2)),
        "insertion": with_insertion_codes(ubq),
        "two_models": two_models(ubq),
        "zero_occupancy": with_zero_occupancy(ubq),
        "ties": cutoff_ties(ubq),
    }
