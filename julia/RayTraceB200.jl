# RayTraceB200.jl — Julia host shim over libraytrace_b200.so (C ABI: include/raytrace_b200.h).
# NOT EXECUTED in this repository's CI: the build image has no Julia.  It shows the binding a maintainer of
# andersontodds/ray-trace adds to switch the hot path over; tests exercise the identical symbols through ctypes.
module RayTraceB200

const LIB = get(ENV, "RAYTRACE_B200_LIB", joinpath(@__DIR__, "..", "ray-trace_b200", "libraytrace_b200.so"))

struct EnvParams      # rt_env_params
    preset::Int32; density_model::Int32
    c::Float64; re::Float64; B0::Float64; e::Float64; me::Float64; mp::Float64; eps::Float64
    Lppi::Float64; Lppo::Float64; ne_Lppi::Float64; ca_coef::Float64; ca_trough::Float64
    iono_n0::Float64; iono_k::Float64; iono_h::Float64
    fd_step_angle::Float64; fd_step_r::Float64   # 0 = the reference's 1e-11 (parity path)
end

mutable struct SolverOpts   # rt_solver_opts
    adaptive::Int32; save_mode::Int32; max_samples::Int32; save_stride::Int32
    maxiters::Int64; t1::Float64; dt::Float64; reltol::Float64; abstol::Float64
    dtmax::Float64; dtmin::Float64; r_esc::Float64; slice_attempts::Int32; coop_threshold::Int32
    single_chain::Int32; nan_policy::Int32
end

struct Result         # rt_result: caller-allocated arrays
    u_end::Ptr{Float64}; t_end::Ptr{Float64}; group_delay::Ptr{Float64}; status::Ptr{Int32}
    naccept::Ptr{Int32}; nreject::Ptr{Int32}; nfe::Ptr{Int32}; nreflect::Ptr{Int32}
end

check(rc, ctx=C_NULL) = rc == 0 || error(unsafe_string(ccall((:rt_last_error, LIB), Cstring, (Ptr{Cvoid},), ctx)))

function create(preset::Integer; device=0)
    ctx = Ref{Ptr{Cvoid}}(C_NULL)
    check(ccall((:rt_create, LIB), Cint, (Ref{Ptr{Cvoid}}, Cint), ctx, device))
    env = Ref{EnvParams}(); check(ccall((:rt_env_preset, LIB), Cint, (Cint, Ref{EnvParams}), preset, env))
    check(ccall((:rt_set_env, LIB), Cint, (Ptr{Cvoid}, Ref{EnvParams}), ctx[], env), ctx[])
    ctx[]
end

# initialize_plasmasphere(Lppi, d, R̄, mlt) -> Lppo          (plasmasphere.jl:48)
function initialize_plasmasphere(Lppi, d, R̄, mlt; re=6.3712e6)
    Lppo = Ref{Float64}(); ne = Ref{Float64}()
    check(ccall((:rt_initialize_plasmasphere, LIB), Cint, (Float64, Float64, Float64, Float64, Float64, Ref{Float64}, Ref{Float64}),
                Lppi, d, R̄, mlt, re, Lppo, ne))
    Lppo[]
end

# phase_refractive_index(r, λ, χ, freq) -> [μ, dμdψ]          (RayTrace_lat.jl:44); vectors in, vectors out
function phase_refractive_index(ctx, r::Vector{Float64}, λ::Vector{Float64}, χ::Vector{Float64}, freq::Vector{Float64})
    n = length(r); μ = similar(r); dμdψ = similar(r)
    check(ccall((:rt_phase_refractive_index_2d, LIB), Cint,
                (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}),
                ctx, n, r, λ, χ, freq, μ, dμdψ), ctx)
    μ, dμdψ
end

# haselgrove!(du, u, p, t) for a batch: u, du are n×4 column-major == [4][n] component-major   (RayTrace_lat.jl:251)
function haselgrove!(du::Matrix{Float64}, u::Matrix{Float64}, p, t; ctx)
    check(ccall((:rt_haselgrove_2d, LIB), Cint, (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}), ctx, size(u, 1), u, du), ctx)
    du
end

# solve(ODEProblem(haselgrove!, u0, tspan, p), Tsit5(); reltol, dtmax, dtmin, callback=re_cb) over a launch list
# (RayTrace_lat.jl:337-338).  launch columns r, θ, χ, f, t as in ray_start.dat (README.md:11).
function solve_2d(ctx, r, θ, χ, f, t0; reltol=1e-7, abstol=1e-6, dtmax=1e6, dtmin=1e-8, tspan_end=5.0e9, maxiters=100_000)
    n = length(r)
    o = SolverOpts(1, 0, 0, 1, maxiters, tspan_end, 0.0, reltol, abstol, dtmax, dtmin, 0.0, 0, 0, 0, 0)
    check(ccall((:rt_set_solver, LIB), Cint, (Ptr{Cvoid}, Ref{SolverOpts}), ctx, o), ctx)
    u_end = Matrix{Float64}(undef, n, 4); t_end = Vector{Float64}(undef, n); T = similar(t_end)
    ints = [Vector{Int32}(undef, n) for _ in 1:5]
    GC.@preserve u_end t_end T ints begin
        res = Result(pointer(u_end), pointer(t_end), pointer(T), pointer.(ints)...)
        check(ccall((:rt_solve_2d, LIB), Cint,
                    (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Ref{Result}),
                    ctx, n, r, θ, χ, f, t0, res), ctx)
    end
    (u=u_end, t=t_end, group_delay=T, retcode=ints[1], naccept=ints[2], nreject=ints[3], nfe=ints[4])
end

destroy(ctx) = ccall((:rt_destroy, LIB), Cvoid, (Ptr{Cvoid},), ctx)
end # module
