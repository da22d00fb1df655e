import sys; sys.path.insert(0,'.')
import arpeggia_b200 as A
for f in ('tests/data/1ubq.pdb','tests/data/6bft.pdb'):
    atoms = A.parse_pdb(f)
    P = A.Packed(atoms, "/", flags=0)
    ctx = A.Context(0)
    for fl in (0,2,1,3):
        r = ctx.contacts_rows(P, 0.1, 6.5, fl)
        print(f, fl, len(r['rows']), r['pairs_screened'], r['pairs_in_cutoff'], r['pairs_classified'])
