"""Deterministic synthetic launch sweeps.  No RNG: plain grids, so every run and every rank
reproduces the same launch list.  Rows follow ray_start.dat: 2D (r, θ, χ, f, t), 3D (r, θ, ϕ, ρ_r, ρ_θ, ρ_ϕ, T0, f, t)."""
import numpy as np

RE = 6.3712e6  # RayTrace_lat.jl:14


def cfg1():
    """The reference's own ray: u0 = [re+1e6, π/4, 0, 1000], RayTrace_lat.jl:333."""
    return np.array([[RE + 1.0e6], [np.pi / 4], [0.0], [1000.0], [0.0]])


def sweep_2d(n_lat=400, n_freq=250, lat_deg=(20.0, 65.0), freq_hz=(1.0e3, 3.0e4), r0=RE + 1.0e6, chi0=0.0):
    """cfg2: launch latitude × frequency (1–30 kHz, geometric), 10^5 rays by default."""
    lat = np.deg2rad(np.linspace(lat_deg[0], lat_deg[1], n_lat))
    f = np.geomspace(freq_hz[0], freq_hz[1], n_freq)
    LAT, F = np.meshgrid(lat, f, indexing="ij")
    n = LAT.size
    return np.ascontiguousarray(np.vstack([np.full(n, r0), LAT.ravel(), np.full(n, chi0), F.ravel(), np.zeros(n)]))


def sweep_3d(n_colat=100, n_lon=100, n_chi=100, colat_deg=(25.0, 70.0), chi_deg=(-60.0, 60.0), alpha=0.1,
             freq=1000.0, r0=RE + 1.0e6):
    """cfg3/cfg5: colatitude × longitude × wave-normal angle χ; ρ = (cosχ, sinχ cosα, sinχ sinα) so that no
    component of ρ is zero (RayTrace_3D.jl:304-310 divides by |ρ_k|); f = 1000 Hz (RayTrace_3D.jl:391)."""
    th = np.deg2rad(np.linspace(colat_deg[0], colat_deg[1], n_colat))
    ph = np.linspace(0.0, 2 * np.pi, n_lon, endpoint=False)
    chi = np.deg2rad(np.linspace(chi_deg[0], chi_deg[1], n_chi))
    TH, PH, CHI = np.meshgrid(th, ph, chi, indexing="ij")
    n = TH.size
    TH, PH, CHI = TH.ravel(), PH.ravel(), CHI.ravel()
    rho = np.vstack([np.cos(CHI), np.sin(CHI) * np.cos(alpha), np.sin(CHI) * np.sin(alpha)])
    return np.ascontiguousarray(
        np.vstack([np.full(n, r0), TH, PH, rho, np.zeros(n), np.full(n, freq), np.zeros(n)]))


def sweep_2d_long(n_lat=1000, n_freq=1000):
    """cfg4: multi-bounce stress sweep, 40°–62° × 1–5 kHz; solve with t1 = 5e10 (scratch.jl:390), maxiters 1e6."""
    return sweep_2d(n_lat, n_freq, (40.0, 62.0), (1.0e3, 5.0e3))


def partition(n, rank, world):
    """Contiguous launch-list partition [lo, hi) of rank `rank` out of `world`."""
    lo = (n * rank) // world
    hi = (n * (rank + 1)) // world
    return lo, hi


def partition_indices(n, rank, world, layout="contiguous"):
    """Launch-list indices of rank `rank`: "contiguous" = [lo, hi) as partition(); "strided" = rank, rank + world, ...;
    "rotated" = strided with the rank order rotated from one block of `world` rays to the next.
    Rays are independent, so any partition gives the same per-ray results; strided balances sweeps whose expensive
    rays cluster in one part of the launch list (cfg3: the long magnetospheric paths sit at high colatitude)."""
    import numpy as _np
    if layout.startswith("blockcyclic"):  # "blockcyclic[:B]": blocks of B rays go round the ranks (block_cyclic_indices)
        return block_cyclic_indices(n, int(layout.split(":")[1]) if ":" in layout else 4096, world, rank)
    if layout.startswith("rootlight"):
        # "rootlight:K" = rotated, except that in every K-th block of `world` rays the ray that would go to rank 0 goes
        # to one of the other ranks (round robin): the gather root traces 1/K fewer rays and spends the time it saves on
        # receiving and placing everybody's records, so that no rank waits for it
        K = int(layout.split(":")[1])
        i = _np.arange(n, dtype=_np.int64)
        b = i // world
        owner = (i % world + b) % world
        if world > 1:
            move = (owner == 0) & (b % K == 0)
            owner[move] = 1 + (b[move] // K) % (world - 1)
        return i[owner == rank]
    if layout == "strided":
        return _np.arange(rank, n, world)
    if layout == "rotated":
        # block b of `world` consecutive rays goes to the ranks rotated by b: rank r takes offset (r - b) mod world.
        # Like strided, but a rank no longer sees only the indices congruent to r modulo gcd(world, inner sweep
        # length): with 100 wave-normal angles innermost and 8 ranks, strided hands each rank 25 of the 100 angles
        # and the per-rank work differs by +-4.5 %.
        b = _np.arange((n + world - 1) // world, dtype=_np.int64)
        idx = b * world + (rank - b) % world
        return idx[idx < n]
    lo, hi = partition(n, rank, world)
    return _np.arange(lo, hi)


def block_cyclic_indices(n, block, world, rank):
    """Launch-list indices of device `rank` under the library's multi-GPU partition (rt_solve_*_multi): blocks of
    `block` consecutive rays go round the `world` devices; local ray j is element (j // block) * block * world +
    rank * block + j % block."""
    import numpy as _np
    idx = _np.arange(n, dtype=_np.int64)
    return idx[(idx // block) % world == rank]


def launch_angle(b, hbi=1.0e6, re=RE):
    """Wave-normal launch angle helper of scratch.jl:725-731: for a ray leaving the ground at angle b from the
    vertical, alpha = acos((re sin^2 b + cos b sqrt((re+hbi)^2 - re^2 sin^2 b)) / (re+hbi)) is the geocentric angle
    travelled up to the launch altitude hbi and chi = b - alpha the wave-normal angle there.  Returns (alpha, chi)."""
    import numpy as _np
    b = _np.asarray(b, dtype=_np.float64)
    alpha = _np.arccos((re * _np.sin(b) ** 2.0 + _np.cos(b) * ((re + hbi) ** 2.0 - re ** 2.0 * _np.sin(b) ** 2.0) ** 0.5)
                       / (re + hbi))
    return alpha, b - alpha


def index_surface(x=9.031e6, y=1.537e6, psi=None):
    """The refractive-index-surface sweep of RayTrace_lat.jl:380-392: at the point (x, y) of the meridian plane,
    lambda = atan|y/x|, r = sqrt(x^2 + y^2), dip = atan(2 tan lambda), and for every wave-normal angle psi
    (default 0:0.001:2pi) the chi that produces it, chi = -psi + 3pi/2 - dip.  Returns (r, lambda, chi[], psi[]);
    feed them to RayTracer.phase_refractive_index(r, lambda, chi, f) — one batch call on the GPU instead of the
    reference's broadcast — for mu(psi) and dmu/dpsi(psi)."""
    import numpy as _np
    if psi is None:
        psi = _np.arange(0.0, 2.0 * _np.pi, 0.001)
    psi = _np.asarray(psi, dtype=_np.float64)
    lam = _np.arctan(abs(y / x))
    r = _np.sqrt(x ** 2 + y ** 2)
    dip = _np.arctan(2.0 * _np.tan(lam))
    chi = -1.0 * psi + 3.0 * _np.pi / 2 - dip
    return float(r), float(lam), chi, psi


def normalize_rho(rt, launch):
    """[EXT] a 3D launch list whose refractive-index vector has the length the medium gives it:
    rho <- rho * mu(r, theta, phi, rho, f) / |rho| (mu depends on rho through its direction only).  The reference
    launches with |rho| = 1 (RayTrace_3D.jl:390), which is not a solution of the dispersion relation; never used
    on the parity path.  `rt` is a RayTracer("ray3d"); one batch call of refractive_index on the GPU."""
    import numpy as _np
    launch = _np.array(launch, dtype=_np.float64, copy=True)
    mu = rt.refractive_index(launch[0], launch[1], launch[2], launch[3], launch[4], launch[5], launch[7])[0]
    norm = _np.sqrt(launch[3] ** 2 + launch[4] ** 2 + launch[5] ** 2)
    launch[3:6] *= mu / norm
    return launch
