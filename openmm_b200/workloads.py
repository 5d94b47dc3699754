"""Synthetic TIP3P water boxes of the atom counts BASELINE.json names, and the system file format shared
with oracle/ref_driver.cpp.

TIP3P parameters: wrappers/python/openmm/app/data/tip3p.xml:20-26 of the reference.  Coordinates: the equilibrated 895-molecule,
30 Angstrom box wrappers/python/openmm/app/data/tip3p.pdb, stored as the fixture tests/golden/tip3p_895.npy (made by
tests/golden/make_fixtures.py), tiled k x k x k, cropped to a cube, molecules clashing across the new periodic boundary removed,
trimmed to the exact count, centres scaled (<~1 %) to 100 atoms/nm^3.  Left-over atoms (N mod 3) are neutral LJ particles.
"""
import os
import struct
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
TIP3P_FIXTURE = os.path.join(os.path.dirname(HERE), "tests", "golden", "tip3p_895.npy")

Q_O, Q_H = -0.834, 0.417
SIG_O, EPS_O = 0.31507524065751241, 0.635968
SIG_H, EPS_H = 1.0, 0.0
M_O, M_H = 15.99943, 1.007947
D_OH, D_HH = 0.09572, 0.15139006545
DUMMY = dict(q=0.0, sigma=0.33, eps=0.4, mass=39.9)

WORKLOADS = {"dhfr": 23558, "apoa1": 92224, "cellulose": 408609, "stmv": 1066628}


def _ideal_water(o, h1, h2):
    """Re-impose the exact rigid TIP3P geometry on PDB-precision coordinates (keeps O, the HOH plane and the bisector)."""
    b = 0.5 * (h1 + h2) - o
    b /= np.linalg.norm(b, axis=1, keepdims=True)
    hh = h2 - h1
    hh -= b * np.sum(hh * b, axis=1, keepdims=True)
    hh /= np.linalg.norm(hh, axis=1, keepdims=True)
    half = 0.5 * D_HH
    along = np.sqrt(D_OH ** 2 - half ** 2)
    return o + along * b - half * hh, o + along * b + half * hh


def water_box(n_atoms, seed=0, density=100.0):
    """Returns a dict describing an n_atoms system: positions [n,3] (float32-representable doubles), box[9], per-atom
    parameters, exclusion pairs, constraints."""
    from scipy.spatial import cKDTree
    rng = np.random.default_rng(seed)
    base = np.load(TIP3P_FIXTURE).astype(np.float64)           # [895, 3 atoms, 3] nm, box 3.0
    L0 = 3.0
    L = (n_atoms / density) ** (1.0 / 3.0)
    nw, extra = divmod(n_atoms, 3)
    k = int(np.ceil(L * 1.06 / L0))
    shifts = np.array([(i, j, l) for i in range(k) for j in range(k) for l in range(k)], dtype=np.float64) * L0
    mols = (base[None, :, :, :] + shifts[:, None, None, :]).reshape(-1, 3, 3)
    # wrap each molecule by its oxygen into the tiled cell so that cropping by O is well defined
    o = mols[:, 0, :]

    def crop(Lc):
        keep = np.all((o >= 0) & (o < Lc), axis=1)
        m = mols[keep]
        # remove molecules that clash with another one across the periodic boundary of the cropped cell
        pts = np.mod(m.reshape(-1, 3), Lc)
        tree = cKDTree(pts, boxsize=Lc)
        pairs = tree.query_pairs(0.135, output_type="ndarray")
        bad = set()
        for a, b in pairs:
            ma, mb = a // 3, b // 3
            if ma != mb and ma not in bad and mb not in bad:
                bad.add(max(ma, mb))
        oo = cKDTree(np.mod(m[:, 0, :], Lc), boxsize=Lc).query_pairs(0.235, output_type="ndarray")
        for a, b in oo:
            if a not in bad and b not in bad:
                bad.add(max(a, b))
        if bad:
            m = np.delete(m, sorted(bad), axis=0)
        return m

    lo, hi = L, L * 1.05
    m = crop(hi)
    while len(m) < nw:
        hi *= 1.02
        m = crop(hi)
    for _ in range(24):
        mid = 0.5 * (lo + hi)
        mm = crop(mid)
        if len(mm) >= nw:
            hi, m = mid, mm
        else:
            lo = mid
    Lc = hi
    if len(m) > nw:
        m = m[np.sort(rng.choice(len(m), nw, replace=False))]
    # scale molecule positions (not their geometry) to the target box
    scale = L / Lc
    o0 = m[:, 0, :].copy()
    m = m - o0[:, None, :] + (o0 * scale)[:, None, :]
    h1, h2 = _ideal_water(m[:, 0, :], m[:, 1, :], m[:, 2, :])
    m[:, 1, :], m[:, 2, :] = h1, h2
    pos = m.reshape(-1, 3)
    if extra:
        tree = cKDTree(np.mod(pos, L), boxsize=L)
        added = []
        while len(added) < extra:
            p = rng.random(3) * L
            d, _ = tree.query(p)
            if d > 0.27 and all(np.linalg.norm(p - a) > 0.5 for a in added):
                added.append(p)
        pos = np.vstack([pos, np.array(added)])
    pos = pos.astype(np.float32).astype(np.float64)            # every platform is fed the float-rounded coordinates
    n = len(pos)
    assert n == n_atoms
    charge = np.empty(n); sigma = np.empty(n); eps = np.empty(n); mass = np.empty(n)
    charge[0:3 * nw:3], charge[1:3 * nw:3], charge[2:3 * nw:3] = Q_O, Q_H, Q_H
    sigma[0:3 * nw:3], sigma[1:3 * nw:3], sigma[2:3 * nw:3] = SIG_O, SIG_H, SIG_H
    eps[0:3 * nw:3], eps[1:3 * nw:3], eps[2:3 * nw:3] = EPS_O, EPS_H, EPS_H
    mass[0:3 * nw:3], mass[1:3 * nw:3], mass[2:3 * nw:3] = M_O, M_H, M_H
    charge[3 * nw:], sigma[3 * nw:], eps[3 * nw:], mass[3 * nw:] = DUMMY["q"], DUMMY["sigma"], DUMMY["eps"], DUMMY["mass"]
    ow = 3 * np.arange(nw, dtype=np.int32)
    excl = np.stack([np.stack([ow, ow + 1], 1), np.stack([ow, ow + 2], 1), np.stack([ow + 1, ow + 2], 1)], 1).reshape(-1, 2)
    con_d = np.tile(np.array([D_OH, D_OH, D_HH]), nw)
    box = np.array([L, 0, 0, 0, L, 0, 0, 0, L], dtype=np.float64)
    # Maxwell-Boltzmann velocities at 300 K, rigid-body consistent enough for a warm-up to fix (constraints are applied on step 1)
    kT = 0.00831446261815324 * 300.0
    vel = rng.normal(size=(nw + extra, 3))
    vmol = np.repeat(vel[:nw] * np.sqrt(kT / (M_O + 2 * M_H)), 3, axis=0)
    vext = vel[nw:] * np.sqrt(kT / DUMMY["mass"])
    vel = np.vstack([vmol, vext]) if extra else vmol
    return dict(n=n, positions=pos, velocities=vel, box=box, charge=charge, sigma=sigma, epsilon=eps, mass=mass,
                exclusions=excl.astype(np.int32), constraints=(excl.astype(np.int32), con_d), num_waters=nw)


def ljpme_water_box(n_atoms):
    """water_box(n_atoms, seed=n_atoms) with CHARMM-style hydrogens, so that the LJPME dispersion grid sees three particle types"""
    sysd = water_box(n_atoms, seed=n_atoms)
    nw = sysd["num_waters"]
    for k in (1, 2):
        sysd["sigma"][k:3 * nw:3] = 0.0400014
        sysd["epsilon"][k:3 * nw:3] = 0.192464
    return sysd


SYSCACHE_DIRS = [os.path.join(os.path.dirname(HERE), ".syscache"), os.path.join(tempfile.gettempdir(), "b200_syscache")]


def cached_water_box(n_atoms, ljpme=False):
    """water_box(n, seed=n) through an on-disk cache: building the million-atom box takes minutes of host time (24 croppings with
    a k-d tree), so the result is kept under <repo>/.syscache (git-ignored; it travels to the GPU box with the snapshot) or the
    temporary directory.  The cache holds exactly what water_box() returned."""
    name = f"waterbox_{n_atoms}.npz"
    sysd = None
    for d in SYSCACHE_DIRS:
        path = os.path.join(d, name)
        if os.path.exists(path):
            try:
                z = np.load(path)
                sysd = {k: z[k] for k in z.files}
                sysd["n"] = int(sysd["n"]); sysd["num_waters"] = int(sysd["num_waters"])
                sysd["positions"] = sysd["positions"].astype(np.float64); sysd["velocities"] = sysd["velocities"].astype(np.float64)
                sysd["constraints"] = (sysd.pop("con_atoms"), sysd.pop("con_dist"))
                break
            except Exception:
                sysd = None
    if sysd is None:
        sysd = water_box(n_atoms, seed=n_atoms)
        for d in SYSCACHE_DIRS:
            try:
                os.makedirs(d, exist_ok=True)
                tmp = os.path.join(d, f".{name}.{os.getpid()}.npz")
                # positions are float32-representable by construction; velocities are kept in full precision
                np.savez_compressed(tmp, con_atoms=sysd["constraints"][0], con_dist=sysd["constraints"][1],
                                    **{k: (v.astype(np.float32) if k == "positions" else v) for k, v in sysd.items() if k != "constraints"})
                os.replace(tmp, os.path.join(d, name))
                break
            except Exception:
                continue
    if ljpme:      # same modification as ljpme_water_box()
        nw = sysd["num_waters"]
        sysd["sigma"] = sysd["sigma"].copy(); sysd["epsilon"] = sysd["epsilon"].copy()
        for k in (1, 2):
            sysd["sigma"][k:3 * nw:3] = 0.0400014
            sysd["epsilon"][k:3 * nw:3] = 0.192464
    return sysd


def make_context(sysd, cutoff=0.9, tol=5e-4, device=0, grid=None, alpha=None, ljpme=False, lj_alpha=None, lj_grid=None, switch_distance=None,
                 method=None, ewald_kmax=None, reaction_field_dielectric=78.3, dispersion_coefficient=0.0, **kw):
    """Create an openmm_b200.Context for a water_box() system with the reference's default PME settings."""
    import openmm_b200 as b
    a, g = b.pme_parameters(cutoff, sysd["box"], tol)
    alpha = a if alpha is None else alpha
    grid = g if grid is None else grid
    ctx = b.Context(sysd["mass"], sysd["box"], constraints=sysd.get("constraints"), device=device, **kw)
    if ljpme and lj_alpha is None:
        lj_alpha, lj_grid = b.pme_parameters(cutoff, sysd["box"], tol, lj=True)
    extra = {} if method is None else dict(method=method)
    ctx.add_nonbonded_force(sysd["charge"], sysd["sigma"], sysd["epsilon"], cutoff, alpha, grid, exclusions=sysd["exclusions"],
                            exceptions=sysd.get("exceptions"), lj_alpha=lj_alpha, lj_grid=lj_grid, switch_distance=switch_distance,
                            ewald_kmax=ewald_kmax, reaction_field_dielectric=reaction_field_dielectric, dispersion_coefficient=dispersion_coefficient, **extra)
    ctx.lj_pme = (lj_alpha, lj_grid)
    ctx.pme = (alpha, grid)
    return ctx


def write_system_file(path, sysd, cutoff, alpha, grid, method=4, dispersion=False, lj_alpha=None, lj_grid=None, switch_distance=None):
    """Binary system description read by oracle/ref_driver.cpp (layout documented there)."""
    n = sysd["n"]
    exc = sysd.get("exceptions")
    with open(path, "wb") as f:
        f.write(b"B2SY")
        version = 3 if switch_distance is not None else (1 if lj_alpha is None else 2)
        f.write(struct.pack("<ii", version, n))
        f.write(np.asarray(sysd["box"], dtype="<f8").tobytes())
        f.write(struct.pack("<dd", cutoff, alpha))
        f.write(struct.pack("<iiii", int(grid[0]), int(grid[1]), int(grid[2]), method))
        f.write(struct.pack("<d", 1.0 if dispersion else 0.0))
        if version >= 2:
            lg = lj_grid if lj_grid is not None else (0, 0, 0)
            f.write(struct.pack("<diii", lj_alpha if lj_alpha is not None else 0.0, int(lg[0]), int(lg[1]), int(lg[2])))
        if version >= 3:
            f.write(struct.pack("<d", switch_distance))
        for key in ("charge", "sigma", "epsilon", "mass"):
            f.write(np.asarray(sysd[key], dtype="<f8").tobytes())
        f.write(np.asarray(sysd["positions"], dtype="<f8").reshape(-1).tobytes())
        vel = sysd.get("velocities")
        f.write(np.asarray(vel if vel is not None else np.zeros((n, 3)), dtype="<f8").reshape(-1).tobytes())
        excl = np.asarray(sysd["exclusions"], dtype=np.int32).reshape(-1, 2)
        rec = np.zeros(len(excl), dtype=[("i", "<i4"), ("j", "<i4"), ("qq", "<f8"), ("sig", "<f8"), ("eps", "<f8")])
        rec["i"], rec["j"], rec["sig"] = excl[:, 0], excl[:, 1], 1.0
        if exc is not None and len(exc[1]) > 0:      # exceptions with parameters: find their rows among the exclusions
            key = excl[:, 0].astype(np.int64) * n + excl[:, 1]
            order = np.argsort(key)
            xpairs = np.asarray(exc[0]).reshape(-1, 2)
            rows = order[np.searchsorted(key[order], xpairs[:, 0].astype(np.int64) * n + xpairs[:, 1])]
            rec["qq"][rows], rec["sig"][rows], rec["eps"][rows] = exc[1], exc[2], exc[3]
        f.write(struct.pack("<i", len(excl)))
        f.write(rec.tobytes())
        con = sysd.get("constraints")
        if con is None or len(con[1]) == 0:
            f.write(struct.pack("<i", 0))
        else:
            ca = np.asarray(con[0], dtype=np.int32).reshape(-1, 2)
            crec = np.zeros(len(ca), dtype=[("i", "<i4"), ("j", "<i4"), ("d", "<f8")])
            crec["i"], crec["j"], crec["d"] = ca[:, 0], ca[:, 1], np.asarray(con[1], dtype=np.float64)
            f.write(struct.pack("<i", len(ca)))
            f.write(crec.tobytes())


def read_eval_file(path, n):
    """Output of `ref_driver eval`: energies, direct and reciprocal forces, optional pair list."""
    with open(path, "rb") as f:
        assert f.read(4) == b"B2EV"
        n2, = struct.unpack("<i", f.read(4))
        assert n2 == n
        e_direct, e_recip = struct.unpack("<dd", f.read(16))
        fd = np.frombuffer(f.read(24 * n), dtype="<f8").reshape(n, 3)
        fr = np.frombuffer(f.read(24 * n), dtype="<f8").reshape(n, 3)
        npairs, = struct.unpack("<q", f.read(8))
        pairs = None
        pair_count = pair_hash = None
        if npairs >= 0:
            pairs = np.frombuffer(f.read(8 * npairs), dtype="<i4").reshape(npairs, 2)
        elif npairs < -1:       # --pair-hash: count and order-independent hash instead of the list
            pair_count = -2 - npairs
            pair_hash, = struct.unpack("<Q", f.read(8))
    return dict(e_direct=e_direct, e_recip=e_recip, f_direct=fd, f_recip=fr, pairs=pairs, pair_count=pair_count, pair_hash=pair_hash)
