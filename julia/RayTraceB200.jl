# RayTraceB200.jl — Julia host shim over libraytrace_b200.so (C ABI: include/raytrace_b200.h).
#
# NOT EXECUTED in this repository: the build image has no Julia.  It is the binding a maintainer of
# andersontodds/ray-trace adds to switch the hot path over.  tests/test_julia_shim.py parses every `ccall` and struct
# in this file and checks names, arity, argument kinds, field order and sizes against the header; the Python mirror
# (ray-trace_b200/api.py) exercises the identical symbols on the GPU.
#
# Array convention: a Julia `Matrix{Float64}(undef, n, dim)` is column-major, i.e. component c of ray i sits at
# i + n*(c-1): exactly the ABI's component-major `[dim][n]`.  So `u[i, :]` is ray i's state, `u[:, 1]` all radii, and
# no transposition happens on either side.  `Bvec` is n×3 the same way.  Trajectory rows are the exception: the ABI
# packs them row-major `[total][columns]`, which a Julia `Matrix(undef, columns, total)` views as columns×total.
module RayTraceB200

const LIB = get(ENV, "RAYTRACE_B200_LIB", joinpath(@__DIR__, "..", "ray-trace_b200", "libraytrace_b200.so"))

const LAT2D, RAY3D, RAYMAIN, SCRATCH = Int32(0), Int32(1), Int32(2), Int32(3)      # rt_preset
const RETCODES = Dict(1 => :Success, 2 => :Terminated, 3 => :Escaped, 4 => :MaxIters,
                      5 => :DtLessThanMin, 6 => :Unstable, 7 => :DtNaN)            # rt_status

mutable struct EnvParams      # rt_env_params
    preset::Int32; density_model::Int32
    c::Float64; re::Float64; B0::Float64; e::Float64; me::Float64; mp::Float64; eps::Float64
    Lppi::Float64; Lppo::Float64; ne_Lppi::Float64; ca_coef::Float64; ca_trough::Float64
    iono_n0::Float64; iono_k::Float64; iono_h::Float64
    fd_step_angle::Float64; fd_step_r::Float64   # 0 = the reference's 1e-11 (parity path)
    deriv_mode::Int32; root_mode::Int32          # [EXT] 1 = analytic d(mu)/dr, d(mu)/dtheta; 1 = root without abs()
    EnvParams() = new()
end

mutable struct SolverOpts   # rt_solver_opts
    adaptive::Int32; save_mode::Int32; max_samples::Int32; save_stride::Int32
    maxiters::Int64; t1::Float64; dt::Float64; reltol::Float64; abstol::Float64
    dtmax::Float64; dtmin::Float64; r_esc::Float64; slice_attempts::Int32; coop_threshold::Int32
    single_chain::Int32; nan_policy::Int32
    SolverOpts() = new()
end

struct Result         # rt_result: caller-allocated arrays
    u_end::Ptr{Float64}; t_end::Ptr{Float64}; group_delay::Ptr{Float64}; status::Ptr{Int32}
    naccept::Ptr{Int32}; nreject::Ptr{Int32}; nfe::Ptr{Int32}; nreflect::Ptr{Int32}
end

mutable struct Stats  # rt_stats
    kernel_launches::Int64; slices::Int64; integrate_ms::Float64; total_ms::Float64
    total_nfe::Int64; prio_rays::Int64
    Stats() = new()
end

lasterror(ctx=C_NULL) = unsafe_string(ccall((:rt_last_error, LIB), Cstring, (Ptr{Cvoid},), ctx))
check(rc, ctx=C_NULL) = rc == 0 || error("raytrace_b200 error $rc: " * lasterror(ctx))

# ---- context -----------------------------------------------------------------------------------------------------
mutable struct Tracer
    ctx::Ptr{Cvoid}
    preset::Int32
    env::EnvParams
    opts::SolverOpts
end
dim(t::Tracer) = t.preset == RAY3D ? 4 + 3 : 4

# create(preset): the script-top constants and the Lppo pre-solve of the named file (RayTrace_lat.jl:13-41,
# RayTrace_3D.jl:38-52) plus that file's solve() keyword arguments (RayTrace_lat.jl:338, RayTrace_3D.jl:395)
function create(preset::Integer; device::Integer=0)
    ctx = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:rt_create, LIB), Cint, (Ref{Ptr{Cvoid}}, Cint), ctx, device))
    env = EnvParams(); opts = SolverOpts()
    ok = false
    try
        check(ccall((:rt_env_preset, LIB), Cint, (Cint, Ref{EnvParams}), preset, env))
        check(ccall((:rt_solver_preset, LIB), Cint, (Cint, Ref{SolverOpts}), preset, opts))
        check(ccall((:rt_set_env, LIB), Cint, (Ptr{Cvoid}, Ref{EnvParams}), ctx[], env), ctx[])
        check(ccall((:rt_set_solver, LIB), Cint, (Ptr{Cvoid}, Ref{SolverOpts}), ctx[], opts), ctx[])
        ok = true
    finally
        ok || ccall((:rt_destroy, LIB), Cvoid, (Ptr{Cvoid},), ctx[])     # no leak when a set-up call fails
    end
    t = Tracer(ctx[], Int32(preset), env, opts)
    finalizer(destroy, t)
    t
end

function destroy(t::Tracer)
    t.ctx == C_NULL || ccall((:rt_destroy, LIB), Cvoid, (Ptr{Cvoid},), t.ctx)
    t.ctx = C_NULL
    nothing
end

version() = ccall((:rt_version, LIB), Cint, ())

function set_env!(t::Tracer, env::EnvParams = t.env)
    check(ccall((:rt_set_env, LIB), Cint, (Ptr{Cvoid}, Ref{EnvParams}), t.ctx, env), t.ctx)
    t.env = env
end

# keyword arguments of solve(): reltol, abstol, dtmax, dtmin, maxiters, tspan end (t1), adaptive/dt, saving
function set_solver!(t::Tracer; kw...)
    for (k, v) in kw
        setfield!(t.opts, k, convert(fieldtype(SolverOpts, k), v))
    end
    check(ccall((:rt_set_solver, LIB), Cint, (Ptr{Cvoid}, Ref{SolverOpts}), t.ctx, t.opts), t.ctx)
    t.opts
end

function stats(t::Tracer)
    s = Stats()
    check(ccall((:rt_get_stats, LIB), Cint, (Ptr{Cvoid}, Ref{Stats}), t.ctx, s), t.ctx)
    s
end

# ---- environment, host side ------------------------------------------------------------------------------------------
# initialize_plasmasphere(Lppi, d, R̄, mlt) -> Lppo          (plasmasphere.jl:48-71, RayTrace_3D.jl:68-91)
function initialize_plasmasphere(Lppi, d, R̄, mlt; re=6.3712e6)
    Lppo = Ref{Float64}(); ne = Ref{Float64}()
    check(ccall((:rt_initialize_plasmasphere, LIB), Cint,
                (Float64, Float64, Float64, Float64, Float64, Ref{Float64}, Ref{Float64}),
                Lppi, d, R̄, mlt, re, Lppo, ne))
    Lppo[]
end

vec64(x, n) = x isa AbstractVector ? convert(Vector{Float64}, x) : fill(Float64(x), n)
nrays(xs...) = maximum(x isa AbstractVector ? length(x) : 1 for x in xs)

# ---- the reference's functions, batched: scalars or vectors in, vectors out ----------------------------------------------
# phase_refractive_index(r, λ, χ, freq) -> [μ, dμdψ]          (RayTrace_lat.jl:44-194; RayMain.jl:125, scratch.jl:63)
function phase_refractive_index(t::Tracer, r, λ, χ, freq)
    n = nrays(r, λ, χ, freq)
    r, λ, χ, freq = vec64(r, n), vec64(λ, n), vec64(χ, n), vec64(freq, n)
    μ = Vector{Float64}(undef, n); dμdψ = Vector{Float64}(undef, n)
    check(ccall((:rt_phase_refractive_index_2d, LIB), Cint,
                (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                t.ctx, n, r, λ, χ, freq, μ, dμdψ), t.ctx)
    μ, dμdψ
end

# refractive_index(r,θ,ϕ,ρ_r,ρ_θ,ρ_ϕ,freq) -> [μ, dμdψ, ψ, Bvec]      (RayTrace_3D.jl:93-219); Bvec is n×3
function refractive_index(t::Tracer, r, θ, ϕ, ρ_r, ρ_θ, ρ_ϕ, freq)
    n = nrays(r, θ, ϕ, ρ_r, ρ_θ, ρ_ϕ, freq)
    a = [vec64(x, n) for x in (r, θ, ϕ, ρ_r, ρ_θ, ρ_ϕ, freq)]
    μ = Vector{Float64}(undef, n); dμdψ = similar(μ); ψ = similar(μ); Bvec = Matrix{Float64}(undef, n, 3)
    check(ccall((:rt_refractive_index_3d, LIB), Cint,
                (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64},
                 Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                t.ctx, n, a[1], a[2], a[3], a[4], a[5], a[6], a[7], μ, dμdψ, ψ, Bvec), t.ctx)
    μ, dμdψ, ψ, Bvec
end

# magnetic_field(r,θ,ϕ) -> [Br, Bθ, Bϕ]      (RayTrace_3D.jl:54-66): the Bvec of refractive_index (any ρ, any freq)
magnetic_field(t::Tracer, r, θ, ϕ) = refractive_index(t, r, θ, ϕ, 1.0, 1.0, 1.0, 1000.0)[4]

# haselgrove!(du,u,p,t) for a batch.  2D: u, du are n×4, p unused (RayTrace_lat.jl:251-296; the frequency is u[:,4]).
# 3D: u, du are n×7 and p = [freq] with freq a scalar or an n-vector (RayTrace_3D.jl:327-359).
function haselgrove!(du::Matrix{Float64}, u::Matrix{Float64}, p, time; tracer::Tracer)
    n = size(u, 1)
    size(u, 2) == dim(tracer) && size(du) == size(u) || error("u and du must be n×$(dim(tracer))")
    if tracer.preset == RAY3D
        freq = vec64(p[1], n)
        check(ccall((:rt_haselgrove_3d, LIB), Cint, (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                    tracer.ctx, n, u, freq, du), tracer.ctx)
    else
        check(ccall((:rt_haselgrove_2d, LIB), Cint, (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}),
                    tracer.ctx, n, u, du), tracer.ctx)
    end
    du
end

# plasmasphere.jl, batched: L_rλ :122, plasmasphere_eq :73-94, diffusive_equilibrium :96-106, ionosphere_eq :108-117,
# ne_total :171 (cm^-3).  2D presets only (the file uses the RayTrace_lat.jl seasonal term).
function plasmasphere(t::Tracer, r, λ)
    n = nrays(r, λ)
    r, λ = vec64(r, n), vec64(λ, n)
    out = Matrix{Float64}(undef, n, 5)
    check(ccall((:rt_plasmasphere_density, LIB), Cint, (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                t.ctx, n, r, λ, out), t.ctx)
    (L = out[:, 1], plasmasphere_eq = out[:, 2], diffusive_equilibrium = out[:, 3], ionosphere_eq = out[:, 4],
     ne_total = out[:, 5])
end
L_rλ(t::Tracer, r, λ) = plasmasphere(t, r, λ).L
plasmasphere_eq(t::Tracer, r, λ) = plasmasphere(t, r, λ).plasmasphere_eq
diffusive_equilibrium(t::Tracer, r, λ) = plasmasphere(t, r, λ).diffusive_equilibrium
ionosphere_eq(t::Tracer, r) = plasmasphere(t, r, 0.0).ionosphere_eq

# the meridian-plane maps of plasmasphere.jl:157-175 and RayTrace_3D.jl:504-517 on x[nx] × y[ny] (metres); each map is
# returned ny×nx (out[iy, ix]), the orientation heatmap!(x, y, z) expects
function meridian_maps(t::Tracer, x::Vector{Float64}, y::Vector{Float64})
    nx, ny = length(x), length(y)
    out = Array{Float64}(undef, ny, nx, 4)
    check(ccall((:rt_meridian_maps, LIB), Cint, (Ptr{Cvoid}, Int64, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                t.ctx, nx, ny, x, y, out), t.ctx)
    (L = out[:, :, 1], ne_total = out[:, :, 2], log10_ne_total = out[:, :, 3], Bmag = out[:, :, 4])
end

# ---- the solve: ODEProblem(haselgrove!, u0, tspan, p) + solve(prob, Tsit5(); ..., callback = re_cb) over a launch list ----
# Result arrays of n rays (Julia owns them; the library writes them through the pointers of `Result`)
struct Terminal
    u::Matrix{Float64}            # n×dim terminal states: sol[end] per ray
    t::Vector{Float64}            # end times
    group_delay::Vector{Float64}  # 2D: u4_end - f0 (RayTrace_lat.jl:257,273); 3D: T (RayTrace_3D.jl:356)
    retcode::Vector{Int32}        # rt_status (RETCODES)
    naccept::Vector{Int32}; nreject::Vector{Int32}; nfe::Vector{Int32}; nreflect::Vector{Int32}
end
Terminal(n, d) = Terminal(Matrix{Float64}(undef, n, d), Vector{Float64}(undef, n), Vector{Float64}(undef, n),
                          (Vector{Int32}(undef, n) for _ in 1:5)...)
result_ptrs(s::Terminal) = Result(pointer(s.u), pointer(s.t), pointer(s.group_delay), pointer(s.retcode),
                                  pointer(s.naccept), pointer(s.nreject), pointer(s.nfe), pointer(s.nreflect))

# 2D launch list: columns r, θ (latitude for LAT2D, colatitude for RAYMAIN / SCRATCH), χ, f, t of ray_start.dat
# (README.md:11).  Replaces RayTrace_lat.jl:333-338; solver options are the Tracer's (set_solver!).
function solve_2d(t::Tracer, r::Vector{Float64}, θ::Vector{Float64}, χ::Vector{Float64}, f::Vector{Float64},
                  t0::Vector{Float64} = zeros(length(r)))
    n = length(r)
    s = Terminal(n, 4)
    GC.@preserve s begin
        res = result_ptrs(s)
        check(ccall((:rt_solve_2d, LIB), Cint,
                    (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ref{Result}),
                    t.ctx, n, r, θ, χ, f, t0, res), t.ctx)
    end
    s
end

# 3D launch list: r, θ, ϕ, ρ_r, ρ_θ, ρ_ϕ, T0, f (= p[1]), t.  Replaces RayTrace_3D.jl:390-395.
function solve_3d(t::Tracer, r::Vector{Float64}, θ::Vector{Float64}, ϕ::Vector{Float64}, ρ_r::Vector{Float64},
                  ρ_θ::Vector{Float64}, ρ_ϕ::Vector{Float64}, T0::Vector{Float64}, f::Vector{Float64},
                  t0::Vector{Float64} = zeros(length(r)))
    n = length(r)
    s = Terminal(n, 7)
    GC.@preserve s begin
        res = result_ptrs(s)
        check(ccall((:rt_solve_3d, LIB), Cint,
                    (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64},
                     Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ref{Result}),
                    t.ctx, n, r, θ, ϕ, ρ_r, ρ_θ, ρ_ϕ, T0, f, t0, res), t.ctx)
    end
    s
end

# ---- trajectories of the last solve (set_solver!(t; save_mode = 1, max_samples = ...)) -------------------------------
# hasel_soln.t, hasel_soln[i,:] and saved_μ (RayTrace_lat.jl:345-349,358-366): rows (t, u..., [μ, dμdψ, dip, ψ]).
# Returns (offsets, rows) with rows columns×total: ray i's samples are rows[:, offsets[i]+1 : offsets[i+1]].
function traj_copy(t::Tracer, n::Integer)
    total = Ref{Int64}(0)
    check(ccall((:rt_traj_size, LIB), Cint, (Ptr{Cvoid}, Ref{Int64}), t.ctx, total), t.ctx)
    ncol = ccall((:rt_traj_columns, LIB), Cint, (Ptr{Cvoid},), t.ctx)
    offsets = Vector{Int64}(undef, n + 1); rows = Matrix{Float64}(undef, ncol, total[])
    check(ccall((:rt_traj_copy, LIB), Cint, (Ptr{Cvoid}, Ptr{Int64}, Ptr{Float64}), t.ctx, offsets, rows), t.ctx)
    offsets, rows
end

# 1 where max_samples was too small for a ray (samples dropped before the final one, which is always kept)
function traj_truncated(t::Tracer, n::Integer)
    flags = Vector{Int32}(undef, n)
    check(ccall((:rt_traj_truncated, LIB), Cint, (Ptr{Cvoid}, Ptr{Int32}), t.ctx, flags), t.ctx)
    flags
end

# x = r.*cos.(λ), y = r.*sin.(λ) (RayTrace_lat.jl:351-352; colatitude files: RayMain.jl:400-401) and the dt series
# (RayTrace_lat.jl:369-372), computed on the device; (x, y, dt) — 3D: (x, y, z, dt) — per sample, columns×total
function traj_derived(t::Tracer, offsets::Vector{Int64})
    nd = ccall((:rt_traj_derived_columns, LIB), Cint, (Ptr{Cvoid},), t.ctx)
    out = Matrix{Float64}(undef, nd, offsets[end])
    check(ccall((:rt_traj_derived, LIB), Cint, (Ptr{Cvoid}, Ptr{Int64}, Ptr{Float64}), t.ctx, offsets, out), t.ctx)
    out
end

# the one-ray call of the reference, as a drop-in: sol.t, sol.u (columns = components), sol.saved_μ, sol.retcode
function solve_ray(t::Tracer, u0::Vector{Float64}, tspan, p = Float64[]; max_samples = 16384)
    set_solver!(t; t1 = tspan[2], save_mode = 1, max_samples = max_samples, save_stride = 1)
    col(x) = [Float64(x)]
    term = t.preset == RAY3D ?
        solve_3d(t, col(u0[1]), col(u0[2]), col(u0[3]), col(u0[4]), col(u0[5]), col(u0[6]), col(u0[7]), col(p[1]), col(tspan[1])) :
        solve_2d(t, col(u0[1]), col(u0[2]), col(u0[3]), col(u0[4]), col(tspan[1]))
    _, rows = traj_copy(t, 1)
    d = dim(t)
    (t = rows[1, :], u = permutedims(rows[2:1+d, :]), saved_μ = permutedims(rows[2+d:end, :]),
     retcode = RETCODES[Int(term.retcode[1])], truncated = traj_truncated(t, 1)[1] != 0, terminal = term)
end

# ---- every GPU of the box from this one Julia process (rt_solve_*_multi) -------------------------------------------------
mutable struct MultiTracer
    m::Ptr{Cvoid}
    preset::Int32
    opts::SolverOpts
end

function multi_create(preset::Integer; ngpu::Integer = 0, devices::Vector{Cint} = Cint[])
    m = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:rt_multi_create, LIB), Cint, (Ref{Ptr{Cvoid}}, Cint, Ptr{Cint}), m,
                isempty(devices) ? ngpu : length(devices), isempty(devices) ? C_NULL : devices))
    env = EnvParams(); opts = SolverOpts()
    ok = false
    try
        check(ccall((:rt_env_preset, LIB), Cint, (Cint, Ref{EnvParams}), preset, env))
        check(ccall((:rt_solver_preset, LIB), Cint, (Cint, Ref{SolverOpts}), preset, opts))
        mcheck(ccall((:rt_multi_set_env, LIB), Cint, (Ptr{Cvoid}, Ref{EnvParams}), m[], env), m[])
        mcheck(ccall((:rt_multi_set_solver, LIB), Cint, (Ptr{Cvoid}, Ref{SolverOpts}), m[], opts), m[])
        ok = true
    finally
        ok || ccall((:rt_multi_destroy, LIB), Cvoid, (Ptr{Cvoid},), m[])
    end
    t = MultiTracer(m[], Int32(preset), opts)
    finalizer(multi_destroy, t)
    t
end
mcheck(rc, m) = rc == 0 || error("raytrace_b200 error $rc: " *
                                 unsafe_string(ccall((:rt_multi_last_error, LIB), Cstring, (Ptr{Cvoid},), m)))
function multi_destroy(t::MultiTracer)
    t.m == C_NULL || ccall((:rt_multi_destroy, LIB), Cvoid, (Ptr{Cvoid},), t.m)
    t.m = C_NULL
    nothing
end
multi_count(t::MultiTracer) = ccall((:rt_multi_count, LIB), Cint, (Ptr{Cvoid},), t.m)

function set_solver!(t::MultiTracer; kw...)
    for (k, v) in kw
        setfield!(t.opts, k, convert(fieldtype(SolverOpts, k), v))
    end
    mcheck(ccall((:rt_multi_set_solver, LIB), Cint, (Ptr{Cvoid}, Ref{SolverOpts}), t.m, t.opts), t.m)
    t.opts
end
set_block!(t::MultiTracer, block::Integer) = mcheck(ccall((:rt_multi_set_block, LIB), Cint, (Ptr{Cvoid}, Int64), t.m, block), t.m)

function solve_2d(t::MultiTracer, r::Vector{Float64}, θ::Vector{Float64}, χ::Vector{Float64}, f::Vector{Float64},
                  t0::Vector{Float64} = zeros(length(r)))
    n = length(r)
    s = Terminal(n, 4)
    GC.@preserve s begin
        res = result_ptrs(s)
        mcheck(ccall((:rt_solve_2d_multi, LIB), Cint,
                     (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ref{Result}),
                     t.m, n, r, θ, χ, f, t0, res), t.m)
    end
    s
end

function solve_3d(t::MultiTracer, r::Vector{Float64}, θ::Vector{Float64}, ϕ::Vector{Float64}, ρ_r::Vector{Float64},
                  ρ_θ::Vector{Float64}, ρ_ϕ::Vector{Float64}, T0::Vector{Float64}, f::Vector{Float64},
                  t0::Vector{Float64} = zeros(length(r)))
    n = length(r)
    s = Terminal(n, 7)
    GC.@preserve s begin
        res = result_ptrs(s)
        mcheck(ccall((:rt_solve_3d_multi, LIB), Cint,
                     (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64},
                      Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ref{Result}),
                     t.m, n, r, θ, ϕ, ρ_r, ρ_θ, ρ_ϕ, T0, f, t0, res), t.m)
    end
    s
end

# which rays device `gpu` (0-based) traced: block length and per-device counts; launch-list index of a local ray
function multi_partition(t::MultiTracer, n::Integer)
    block = Ref{Int64}(0); counts = Vector{Int64}(undef, multi_count(t))
    mcheck(ccall((:rt_multi_partition, LIB), Cint, (Ptr{Cvoid}, Int64, Ref{Int64}, Ptr{Int64}), t.m, n, block, counts), t.m)
    block[], counts
end
multi_global_index(t::MultiTracer, n::Integer, gpu::Integer, loc::Integer) =
    ccall((:rt_multi_global_index, LIB), Int64, (Ptr{Cvoid}, Int64, Cint, Int64), t.m, n, gpu, loc)
function multi_stats(t::MultiTracer, gpu::Integer)
    s = Stats()
    mcheck(ccall((:rt_multi_get_stats, LIB), Cint, (Ptr{Cvoid}, Cint, Ref{Stats}), t.m, gpu, s), t.m)
    s
end

# ---- ray_start.dat (README.md:11): r θ χ f t per line (3D: r θ ϕ ρ_r ρ_θ ρ_ϕ T0 f t), '#' comments ---------------------
function read_ray_start(path::AbstractString)
    rows = Vector{Vector{Float64}}()
    for line in eachline(path)
        s = strip(split(line, '#')[1])
        isempty(s) || push!(rows, parse.(Float64, split(s)))
    end
    isempty(rows) && return Matrix{Float64}(undef, 0, 5)
    permutedims(reduce(hcat, rows))     # n×5 or n×9: column k is the k-th launch column
end

end # module
