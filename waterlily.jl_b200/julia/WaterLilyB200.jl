# WaterLilyB200.jl -- thin `ccall` shim that plugs libwaterlily_b200.so into WaterLily's own
# `Simulation` through the documented factory hooks `flow_ctor` / `pois_ctor`
# (reference src/WaterLily.jl:69-74,93-106).  No KernelAbstractions, no CUDA.jl, no kernels here:
# every numerical operation is a single call into the C ABI of include/wl_b200.h.
#
# STATUS: Julia is not installed in the build image or on the GPU box, so this file has been
# syntax-reviewed only; the same ABI is exercised by the ctypes mirror in ../__init__.py, whose
# tests are the executable specification of every call made below.
#
#   using WaterLily, WaterLilyB200
#   sim = B200Simulation((512,256,256), (1,0,0), 64; ν=64/3700, body=B200Sphere((127,127,127),32))
#   sim_step!(sim, 10.0; remeasure=false)          # WaterLily's own time loop drives the library
#   pressure_force(sim); Array(sim.flow.u); sim.pois.n; sim.flow.Δt
module WaterLilyB200

using WaterLily
import WaterLily: AbstractFlow, AbstractPoisson, AbstractBody, mom_step!, measure!, update!, solver!, CFL,
                  pressure_force, viscous_force, total_force, pressure_moment, viscous_moment, total_moment, time

const LIB = get(ENV, "WATERLILY_B200_LIB", joinpath(@__DIR__, "..", "libwaterlily_b200.so"))

# ---- POD mirrors of include/wl_b200.h ------------------------------------------------------------
struct WLConfig
    D::Int32; N::NTuple{3,Int32}; perdir::NTuple{3,Int32}; exitBC::Int32; scheme::Int32
    nu::Float32; dt0::Float32; uBC::NTuple{3,Float32}; maxlevels::Int32; tol::Float32; itmx::Int32
    device::Int32; flags::Int32
end
struct WLBody
    kind::Int32; axis::Int32; center::NTuple{3,Float32}; radius::Float32; minor::Float32
    vel::NTuple{3,Float32}; acc::NTuple{3,Float32}; amp::NTuple{3,Float32}; omega::Float32
end
struct WLNode                                   # wl_body_node (include/wl_b200.h)
    op::Int32; kind::Int32; axis::Int32; radius::Float32; minor::Float32
    x0::NTuple{3,Float32}; theta::NTuple{3,Float32}; xp::NTuple{3,Float32}; V::NTuple{3,Float32}; omega::NTuple{3,Float32}
end
struct WLStepInfo
    dt_next::Float32; n_pred::Int32; n_corr::Int32; flags::Int32; trace_len::NTuple{2,Int32}
    rinf::NTuple{68,Float32}; r1::NTuple{68,Float32}; omega::NTuple{68,Float32}
end

const FIELD = (u=0, u⁰=1, f=2, p=3, σ=4, V=5, μ₀=6, μ₁=7, x=10, r=11, ϵ=12, z=13, L=14, D=15, iD=16)
pad3(t, z) = ntuple(i -> i <= length(t) ? t[i] : z, 3)

mutable struct Handle
    ptr::Ptr{Cvoid}
    function Handle(p)
        h = new(p)
        finalizer(h -> (h.ptr == C_NULL || ccall((:wl_destroy, LIB), Cvoid, (Ptr{Cvoid},), h.ptr); h.ptr = C_NULL), h)
    end
end
lasterr(h) = unsafe_string(ccall((:wl_last_error, LIB), Cstring, (Ptr{Cvoid},), h))
check(h::Handle, rc) = rc == 0 ? nothing : error("libwaterlily_b200 ($rc): ", lasterr(h.ptr))

# ---- device array handle: `Array(a)` downloads, `copyto!(a, ::Array)` uploads ---------------------
struct B200Array{T,N} <: AbstractArray{T,N}
    h::Handle; field::Int32; level::Int32; dims::NTuple{N,Int}
end
Base.size(a::B200Array) = a.dims
function Base.Array(a::B200Array{T,N}) where {T,N}
    out = Array{T,N}(undef, a.dims)
    check(a.h, ccall((:wl_download, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{Float32}, Csize_t),
                     a.h.ptr, a.field, a.level, out, length(out)))
    out
end
function Base.copyto!(a::B200Array{T,N}, src::Array{T,N}) where {T,N}
    check(a.h, ccall((:wl_upload, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ptr{Float32}, Csize_t),
                     a.h.ptr, a.field, a.level, src, length(src)))
    a
end
Base.getindex(a::B200Array, I...) = Array(a)[I...]        # slow scalar probe, tests only
Base.maximum(a::B200Array) = maximum(Array(a))

# ---- parametric bodies the device-side measure! understands ---------------------------------------
struct B200Body <: AbstractBody
    c::WLBody
end
B200Sphere(center, radius; vel=(0,0,0), acc=(0,0,0), amp=(0,0,0), ω=0) =
    B200Body(WLBody(1, 0, pad3(Float32.(center), 0f0), radius, 0, pad3(Float32.(vel), 0f0), pad3(Float32.(acc), 0f0), pad3(Float32.(amp), 0f0), ω))
B200Cylinder(center, radius, axis; kw...) = (b = B200Sphere(center, radius; kw...).c;
    B200Body(WLBody(2, axis - 1, b.center, b.radius, 0, b.vel, b.acc, b.amp, b.omega)))
B200Torus(center, R, r, axis; kw...) = (b = B200Sphere(center, R; kw...).c;
    B200Body(WLBody(3, axis - 1, b.center, R, r, b.vel, b.acc, b.amp, b.omega)))

# ---- library sdfs under a `RigidMap`, combined with WaterLily's own lazy `SetBody` algebra (+ ∪ ∩ -) -----------
# `body.map` is a real `WaterLily.RigidMap`, so `setmap(body; θ, x₀, ...)` (src/RigidMap.jl:54-56) works unchanged,
# on single leaves and through `SetBody` trees alike.
struct B200Rigid{M<:WaterLily.RigidMap} <: AbstractBody
    kind::Int32; axis::Int32; radius::Float32; minor::Float32
    map::M
end
B200Rigid(kind::Symbol, map::WaterLily.RigidMap; radius=0, minor=0, axis=3) =
    B200Rigid(Int32((sphere=1, circle=1, cylinder=2, torus=3, capsule=4)[kind]), Int32(axis - 1), Float32(radius), Float32(minor), map)
rot3(θ::Number) = (0f0, 0f0, Float32(θ)); rot3(θ) = pad3(Float32.(θ), 0f0)       # 2D: θ and ω turn about z
node(b::B200Rigid) = (m = b.map; WLNode(0, b.kind, b.axis, b.radius, b.minor, pad3(Float32.(m.x₀), 0f0), rot3(m.θ),
                                        pad3(Float32.(m.xₚ), 0f0), pad3(Float32.(m.V), 0f0), rot3(m.ω)))
const OPNODE = Dict(1 => WLNode(1, 0, 0, 0, 0, ntuple(_ -> ntuple(_ -> 0f0, 3), 5)...),
                    2 => WLNode(2, 0, 0, 0, 0, ntuple(_ -> ntuple(_ -> 0f0, 3), 5)...),
                    3 => WLNode(3, 0, 0, 0, 0, ntuple(_ -> ntuple(_ -> 0f0, 3), 5)...))
"Postfix program of a body, or `nothing` when a leaf is not a library primitive (then the host path measures it)."
nodes(b::B200Rigid) = WLNode[node(b)]
nodes(::AbstractBody) = nothing
nodes(b::WaterLily.SetBody{typeof(-)}) = (a = nodes(b.a); a === nothing ? nothing : [a; OPNODE[3]])
function nodes(b::WaterLily.SetBody)
    op = b.op === min ? 1 : b.op === max ? 2 : 0
    a, c = nodes(b.a), nodes(b.b)
    (op == 0 || a === nothing || c === nothing || length(a) + length(c) + 1 > 8) ? nothing : [a; c; OPNODE[op]]
end
const CsgBody = Union{B200Rigid,WaterLily.SetBody}

# ---- Flow ------------------------------------------------------------------------------------------
struct B200Flow{D,T} <: AbstractFlow{D,T}
    h::Handle
    u::B200Array{T}; u⁰::B200Array{T}; f::B200Array{T}; p::B200Array{T}; σ::B200Array{T}
    V::B200Array{T}; μ₀::B200Array{T}; μ₁::B200Array{T}
    uBC::Union{NTuple{D,T},Function}; Δt::Vector{T}; ν::T; g::Union{Function,Nothing}; exitBC::Bool; perdir::NTuple
    koff::Int                                        # global index of the slab's first plane (0 on one GPU)
end
const SCHEME = Dict(WaterLily.quick => 0, WaterLily.cds => 1, WaterLily.vanLeer => 2)

# ---- multi-GPU: one Julia process per GPU (e.g. MPI.jl or Distributed), slabs along the last dimension.
# `comm = (rank, nranks, id)`; rank 0 obtains `id = nccl_unique_id()` and ships the 128 bytes to the
# other ranks with whatever transport the host program uses (MPI.Bcast!, a file, a socket).
struct WLDist; rank::Int32; nranks::Int32; id::NTuple{128,UInt8}; end
function nccl_unique_id()
    id = Ref(ntuple(_ -> 0x00, 128))
    ccall((:wl_nccl_unique_id, LIB), Cint, (Ref{NTuple{128,UInt8}},), id) == 0 || error("wl_nccl_unique_id: ", lasterr(C_NULL))
    id[]
end

"`flow_ctor` for `Simulation` (reference src/Flow.jl:133-147).  With `comm`, `N` is the GLOBAL grid and the
returned flow holds this rank's slab: arrays have `N[end]÷nranks+2` planes and `loc` is offset by the slab origin."
function B200Flow(N::NTuple{D}, uBC; Δt=0.25, ν=0., g=nothing, u0=nothing, uλ=nothing, perdir=(), exitBC=false,
                  λ=WaterLily.quick, T=Float32, mem=Array, device=0, comm=nothing) where D
    T == Float32 || error("WL_ERR_UNSUPPORTED: the B200 path is Float32 only")
    haskey(SCHEME, λ) || error("WL_ERR_UNSUPPORTED: λ must be quick, cds or vanLeer")
    # A Function uBC(i,x,t) / g(i,x,t) stays on the host: it is SAMPLED into the boundary tables / acceleration fields
    # the device path consumes (set_bc_function!, set_accel_field!) -- at t = 0 here, at t₁ / (t₀,t₁) in every mom_step!.
    ubc0 = uBC isa Function ? ntuple(_ -> 0f0, D) : Float32.(uBC)
    cfg = Ref(WLConfig(D, pad3(Int32.(N), Int32(0)), ntuple(i -> Int32(i in perdir), 3), exitBC, SCHEME[λ], ν, Δt,
                       pad3(ubc0, 0f0), 10, 2f-3, 32, device, 0))
    out = Ref{Ptr{Cvoid}}(C_NULL)
    koff = 0
    if isnothing(comm) || comm[2] == 1
        rc = ccall((:wl_create, LIB), Cint, (Ref{WLConfig}, Ref{Ptr{Cvoid}}), cfg, out)
    else
        rank, nranks, id = comm
        rc = ccall((:wl_create_slab, LIB), Cint, (Ref{WLConfig}, Ref{WLDist}, Ref{Ptr{Cvoid}}),
                   cfg, Ref(WLDist(rank, nranks, id)), out)
        koff = rank * (N[end] ÷ nranks)
        N = (N[1:end-1]..., N[end] ÷ nranks)           # everything below is the local slab
    end
    rc == 0 || error("wl_create ($rc): ", lasterr(C_NULL))
    h = Handle(out[]); Ng = N .+ 2
    arr(f, dims) = B200Array{T,length(dims)}(h, Int32(f), Int32(0), dims)
    flow = B200Flow{D,T}(h, arr(FIELD.u, (Ng..., D)), arr(FIELD.u⁰, (Ng..., D)), arr(FIELD.f, (Ng..., D)),
                         arr(FIELD.p, Ng), arr(FIELD.σ, Ng), arr(FIELD.V, (Ng..., D)), arr(FIELD.μ₀, (Ng..., D)),
                         arr(FIELD.μ₁, (Ng..., D, D)), uBC isa Function ? uBC : T.(uBC), T[Δt], T(ν), g, exitBC, perdir, koff)
    u0 = isnothing(u0) ? uλ : u0
    if uBC isa Function
        set_bc_function!(flow, uBC; t=zero(T))         # BC!(u,uBC,exitBC,perdir) of the constructor, t = 0 (Flow.jl:141)
        isnothing(u0) && (u0 = (i, x) -> uBC(i, x, zero(T)))               # ic_function(uBC)
    end
    if !isnothing(u0)                                  # apply!(u0,u) on the host, then BC!/exitBC!/u⁰ on the device
        host = Array{T}(undef, Ng..., D)
        shift(x) = x .+ ntuple(d -> d == D ? T(koff) : zero(T), D)        # slab origin in the global grid
        u0 isa Function ? WaterLily.apply!((i, x) -> u0(i, shift(x)), host) : WaterLily.apply!((i, x) -> u0[i], host)
        copyto!(flow.u, host)
        check(h, ccall((:wl_finish_ic, LIB), Cint, (Ptr{Cvoid}, Ref{NTuple{3,Float32}}), h.ptr, Ref(pad3(ubc0, 0f0))))
    end
    flow
end

# ---- Poisson -----------------------------------------------------------------------------------------
struct B200Level{T}
    x::B200Array{T}; r::B200Array{T}; ϵ::B200Array{T}; z::B200Array{T}; L::B200Array{T}; D::B200Array{T}; iD::B200Array{T}
end
struct B200Poisson{T} <: AbstractPoisson{T,B200Array{T},B200Array{T}}
    h::Handle; x::B200Array{T}; L::B200Array{T}; z::B200Array{T}; levels::Vector{B200Level{T}}; n::Vector{Int16}; perdir::NTuple
end
"`pois_ctor` for `Simulation` (reference src/WaterLily.jl:97, src/MultiLevelPoisson.jl:68-76)."
function B200Poisson(flow::B200Flow{D,T}) where {D,T}
    h = flow.h
    rc = ccall((:wl_build_poisson, LIB), Cint, (Ptr{Cvoid},), h.ptr)
    rc == 0 || (occursin("size=a2", lasterr(h.ptr)) ? throw(AssertionError("MultiLevelPoisson requires size=a2ⁿ, where n>2")) : check(h, rc))
    nl = ccall((:wl_nlevels, LIB), Cint, (Ptr{Cvoid},), h.ptr)
    levels = map(0:nl-1) do l
        dims = Ref(ntuple(_ -> Int32(0), 5)); nd = Ref{Int32}(0)
        ccall((:wl_field_dims, LIB), Cint, (Ptr{Cvoid}, Cint, Cint, Ref{NTuple{5,Int32}}, Ref{Int32}), h.ptr, FIELD.x, l, dims, nd)
        ng = Int.(dims[][1:D]); a(f, d) = B200Array{T,length(d)}(h, Int32(f), Int32(l), d)
        B200Level{T}(a(FIELD.x, ng), a(FIELD.r, ng), a(FIELD.ϵ, ng), a(FIELD.z, ng), a(FIELD.L, (ng..., D)), a(FIELD.D, ng), a(FIELD.iD, ng))
    end
    B200Poisson{T}(h, flow.p, flow.μ₀, flow.σ, levels, Int16[], flow.perdir)
end

# ---- the generic functions WaterLily's time loop calls (src/WaterLily.jl:128-149) --------------------
function mom_step!(a::B200Flow{D,T}, b::B200Poisson; udf=nothing, kwargs...) where {D,T}
    isnothing(udf) || error("WL_ERR_UNSUPPORTED: udf hooks cannot cross the C ABI")
    ccall((:wl_set_dt, LIB), Cint, (Ptr{Cvoid}, Cfloat), a.h.ptr, a.Δt[end])
    t₁ = sum(a.Δt); t₀ = t₁ - a.Δt[end]                 # Flow.jl:157
    if a.uBC isa Function                               # all four BC! of the step use t₁ (Flow.jl:194,209,231)
        set_bc_function!(a, a.uBC; t=t₁)
    end
    if a.g isa Function || a.uBC isa Function           # accelerate!(r,t,g,uBC): g + ∂ₜuBC by AD, as Flow.jl:69-73
        acc(i, x, t) = (a.g isa Function ? a.g(i, x, t) : zero(T)) +
                       (a.uBC isa Function ? WaterLily.derivative(τ -> a.uBC(i, x, τ), t) : zero(T))
        set_accel_field!(a, 0, acc, t₀); set_accel_field!(a, 1, acc, t₁)
    end
    info = Ref{WLStepInfo}()
    check(a.h, ccall((:wl_mom_step, LIB), Cint, (Ptr{Cvoid}, Ptr{Cfloat}, Ptr{Cfloat}, Ptr{Cfloat}, Ref{WLStepInfo}),
                     a.h.ptr, C_NULL, C_NULL, C_NULL, info))
    # the solver trace the reference writes with @log (Flow.jl:159,163, MultiLevelPoisson.jl:112,117): same lines, same order
    for (q, tag) in enumerate(("p", "c"))
        WaterLily.@log tag
        for it in 0:info[].trace_len[q]-1
            k = 34 * (q - 1) + it + 1
            WaterLily.@log ", $it, $(info[].rinf[k]), $(info[].r1[k]), $(info[].omega[k])\n"
        end
    end
    push!(b.n, info[].n_pred, info[].n_corr)            # MultiLevelPoisson.jl:127
    push!(a.Δt, info[].dt_next)                         # Flow.jl:166
end
measure!(a::B200Flow, ::WaterLily.NoBody; kw...) = nothing
function measure!(a::B200Flow{D,T}, body::B200Body; t=zero(T), ϵ=1) where {D,T}
    check(a.h, ccall((:wl_measure_body, LIB), Cint, (Ptr{Cvoid}, Ref{WLBody}, Cdouble, Cdouble), a.h.ptr, Ref(body.c), t, ϵ))
end
"Arbitrary `AutoBody` closures: measured by the reference itself on a host `Flow`, then uploaded.
`SetBody` trees of `B200Rigid` leaves go to the device instead (`wl_measure_csg`)."
function measure!(a::B200Flow{D,T}, body::AbstractBody; t=zero(T), ϵ=1) where {D,T}
    nd = nodes(body)
    if nd !== nothing
        check(a.h, ccall((:wl_measure_csg, LIB), Cint, (Ptr{Cvoid}, Ptr{WLNode}, Cint, Cdouble), a.h.ptr, nd, length(nd), ϵ))
        return
    end
    host = WaterLily.Flow(size(a.p) .- 2, ntuple(_ -> zero(T), D); T, perdir=a.perdir, exitBC=a.exitBC)
    fill!(host.V, 0); fill!(host.μ₀, 1); fill!(host.μ₁, 0)
    WaterLily.measure!(host, shifted(body, a.koff, Val(D)); t, ϵ)   # includes the two BC! calls; the library re-applies them (idempotent)
    check(a.h, ccall((:wl_set_body_coeffs, LIB), Cint, (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}),
                     a.h.ptr, host.μ₀, host.μ₁, host.V, host.σ))
end
"A body seen from a slab whose first plane is global plane `koff`: the same body with `x[D] += koff`."
struct ShiftedBody{B<:AbstractBody,D} <: AbstractBody; body::B; koff::Int; end
shifted(b::AbstractBody, koff, ::Val{D}) where D = koff == 0 ? b : ShiftedBody{typeof(b),D}(b, koff)
WaterLily.measure(s::ShiftedBody{B,D}, x, t; kw...) where {B,D} =
    WaterLily.measure(s.body, x .+ ntuple(d -> d == D ? eltype(x)(s.koff) : zero(eltype(x)), D), t; kw...)
WaterLily.sdf(s::ShiftedBody{B,D}, x, t=0; kw...) where {B,D} =
    WaterLily.sdf(s.body, x .+ ntuple(d -> d == D ? eltype(x)(s.koff) : zero(eltype(x)), D), t; kw...)

"""
`pressure_force(flow, body)` for ANY `AbstractBody` (src/Metrics.jl:116-133): `nds(body, loc(0,I), t)` is evaluated by the
reference's own code on the host over the interior cells, the non-zero entries travel through `wl_set_force_weights`, and
the device forms `p[I]·nds` in Float32 and sums in Float64 as the reference does.  Library bodies (`B200Body`, `B200Rigid`
trees) have device-side methods below and never come here.
"""
function pressure_force(a::B200Flow{D,T}, body::AbstractBody) where {D,T}
    Ng = size(a.p); b = shifted(body, a.koff, Val(D)); t = time(a)
    idx = Int64[]; w = T[]
    lin = LinearIndices(Ng)
    for I in WaterLily.inside(a.p)                      # `inside` takes the array (src/core.jl:47); B200Array has `size`
        n = WaterLily.nds(b, WaterLily.loc(0, I, T), t)
        iszero(n) && continue
        push!(idx, lin[I] - 1); append!(w, n)           # 0-based, first dimension fastest; D weights per entry
    end
    check(a.h, ccall((:wl_set_force_weights, LIB), Cint, (Ptr{Cvoid}, Int64, Ptr{Int64}, Ptr{Cfloat}), a.h.ptr, length(idx), idx, w))
    out = Ref(ntuple(_ -> 0.0, 3))
    check(a.h, ccall((:wl_pressure_force_weights, LIB), Cint, (Ptr{Cvoid}, Ref{NTuple{3,Float64}}), a.h.ptr, out))
    collect(out[][1:D])
end
update!(b::B200Poisson) = check(b.h, ccall((:wl_update_poisson, LIB), Cint, (Ptr{Cvoid},), b.h.ptr))
function solver!(b::B200Poisson; kw...)
    n = Ref{Int32}(0)
    check(b.h, ccall((:wl_solver, LIB), Cint, (Ptr{Cvoid}, Ref{Int32}, Ptr{Int32}, Ptr{Cfloat}, Ptr{Cfloat}, Ptr{Cfloat}),
                     b.h.ptr, n, C_NULL, C_NULL, C_NULL, C_NULL))
    push!(b.n, n[])
end
"""
Single-level counterpart of `Poisson(x,L,z)` (reference src/Poisson.jl:22-39): the same device fields, but
`solver!` runs the Jacobi-preconditioned `pcg!` loop of src/Poisson.jl:166-186,212-223 (`wl_poisson_solver`).
Stand-alone solves only (`Poisson_setup`-style use, test/test_poisson.jl:1-26); `mom_step!` always projects
with the multigrid solver.
"""
struct B200PCGPoisson{T} <: AbstractPoisson{T,B200Array{T},B200Array{T}}
    ml::B200Poisson{T}
end
B200PCGPoisson(flow::B200Flow) = B200PCGPoisson(B200Poisson(flow))
Base.getproperty(p::B200PCGPoisson, s::Symbol) = s === :ml ? getfield(p, :ml) : getproperty(getfield(p, :ml), s)
update!(p::B200PCGPoisson) = update!(p.ml)
function solver!(p::B200PCGPoisson; itmx=1e3, kw...)
    n = Ref{Int32}(0)
    check(p.h, ccall((:wl_poisson_solver, LIB), Cint, (Ptr{Cvoid}, Int32, Ref{Int32}), p.h.ptr, Int32(min(itmx, typemax(Int32))), n))
    push!(p.n, n[])
end
function CFL(a::B200Flow; Δt_max=10)
    dt = Ref{Cfloat}(0)
    check(a.h, ccall((:wl_cfl, LIB), Cint, (Ptr{Cvoid}, Ref{Cfloat}), a.h.ptr, dt)); min(Δt_max, dt[])
end
function pressure_force(a::B200Flow{D}, body::B200Body) where D
    out = Ref(ntuple(_ -> 0.0, 3))
    check(a.h, ccall((:wl_pressure_force, LIB), Cint, (Ptr{Cvoid}, Ref{WLBody}, Cdouble, Ref{NTuple{3,Float64}}),
                     a.h.ptr, Ref(body.c), time(a), out))
    collect(out[][1:D])
end

function viscous_force(a::B200Flow{D}, body::B200Body) where D
    out = Ref(ntuple(_ -> 0.0, 3))
    check(a.h, ccall((:wl_viscous_force, LIB), Cint, (Ptr{Cvoid}, Ref{WLBody}, Cdouble, Ref{NTuple{3,Float64}}),
                     a.h.ptr, Ref(body.c), time(a), out))
    collect(out[][1:D])
end
total_force(a::B200Flow, body::B200Body) = pressure_force(a, body) .+ viscous_force(a, body)
function force_moment(x₀, a::B200Flow{D}, body::B200Body, viscous::Bool) where D
    out = Ref(ntuple(_ -> 0.0, 3)); x = Ref(ntuple(i -> i <= D ? Float64(x₀[i]) : 0.0, 3))
    check(a.h, ccall((:wl_force_moment, LIB), Cint, (Ptr{Cvoid}, Ref{WLBody}, Cdouble, Ref{NTuple{3,Float64}}, Cint, Ref{NTuple{3,Float64}}),
                     a.h.ptr, Ref(body.c), time(a), x, viscous, out))
    collect(out[][1:D])
end
pressure_moment(x₀, a::B200Flow, body::B200Body) = force_moment(x₀, a, body, false)
viscous_moment(x₀, a::B200Flow, body::B200Body) = force_moment(x₀, a, body, true)
total_moment(x₀, a::B200Flow, body::B200Body) = pressure_moment(x₀, a, body) .+ viscous_moment(x₀, a, body)

function csg_call(sym, a::B200Flow{D}, body, extra...) where D
    nd = nodes(body)
    if nd === nothing                                     # a tree with a non-library leaf: the host-weights route
        sym === :wl_pressure_force_csg && return invoke(pressure_force, Tuple{B200Flow,AbstractBody}, a, body)
        error("WL_ERR_UNSUPPORTED: viscous force and moments need library primitives (B200Rigid leaves)")
    end
    out = Ref(ntuple(_ -> 0.0, 3))
    sym === :wl_force_moment_csg ?
        check(a.h, ccall((:wl_force_moment_csg, LIB), Cint, (Ptr{Cvoid}, Ptr{WLNode}, Cint, Ref{NTuple{3,Float64}}, Cint, Ref{NTuple{3,Float64}}),
                         a.h.ptr, nd, length(nd), Ref(ntuple(i -> i <= D ? Float64(extra[1][i]) : 0.0, 3)), extra[2], out)) :
    sym === :wl_viscous_force_csg ?
        check(a.h, ccall((:wl_viscous_force_csg, LIB), Cint, (Ptr{Cvoid}, Ptr{WLNode}, Cint, Ref{NTuple{3,Float64}}), a.h.ptr, nd, length(nd), out)) :
        check(a.h, ccall((:wl_pressure_force_csg, LIB), Cint, (Ptr{Cvoid}, Ptr{WLNode}, Cint, Ref{NTuple{3,Float64}}), a.h.ptr, nd, length(nd), out))
    collect(out[][1:D])
end
pressure_force(a::B200Flow, body::CsgBody) = csg_call(:wl_pressure_force_csg, a, body)
viscous_force(a::B200Flow, body::CsgBody) = csg_call(:wl_viscous_force_csg, a, body)
total_force(a::B200Flow, body::CsgBody) = pressure_force(a, body) .+ viscous_force(a, body)
pressure_moment(x₀, a::B200Flow, body::CsgBody) = csg_call(:wl_force_moment_csg, a, body, x₀, 0)
viscous_moment(x₀, a::B200Flow, body::CsgBody) = csg_call(:wl_force_moment_csg, a, body, x₀, 1)
total_moment(x₀, a::B200Flow, body::CsgBody) = pressure_moment(x₀, a, body) .+ viscous_moment(x₀, a, body)

# ---- Function uBC(i,x,t): sampled where BC! evaluates it (src/core.jl:201-219) and handed over as tables ------------
# Layout of wl_set_bc_tables: component outer, direction inner, three planes (index 1, 2, Ng[j]) over the other dims in
# column-major order; normal component: the Dirichlet values; tangential: uBC(I)-uBC(I±δ) on the two ghost planes.
# (The ctypes mirror's `bctables.py` is the executable twin of this function: tests/test_bctables_cpu.py, tests/test_fnbc_gpu.py.)
function bc_tables(uBC::Function, Ng::NTuple{D,Int}, t, perdir, ::Type{T}) where {D,T}
    tab = T[]
    for i in 1:D, j in 1:D, (s, k) in enumerate((1, 2, Ng[j])), I in WaterLily.slice(Ng, k, j)
        v = if j in perdir || (i != j && s == 2)
            zero(T)
        elseif i == j
            T(uBC(i, WaterLily.loc(i, I, T), t))
        else
            J = s == 1 ? I + WaterLily.δ(j, I) : I - WaterLily.δ(j, I)
            T(uBC(i, WaterLily.loc(i, I, T), t)) - T(uBC(i, WaterLily.loc(i, J, T), t))
        end
        push!(tab, v)
    end
    tab
end
"A Function of (i,x,t) seen from a slab whose first plane is global plane `koff`: local indices, global coordinates."
slab_fn(f::Function, koff, ::Val{D}, ::Type{T}) where {D,T} =
    koff == 0 ? f : (i, x, t) -> f(i, x .+ ntuple(d -> d == D ? T(koff) : zero(T), D), t)
function set_bc_function!(a::B200Flow{D,T}, uBC::Function; t=zero(T)) where {D,T}
    tab = bc_tables(slab_fn(uBC, a.koff, Val(D), T), size(a.p), T(t), a.perdir, T)
    check(a.h, ccall((:wl_set_bc_tables, LIB), Cint, (Ptr{Cvoid}, Ptr{Cfloat}, Csize_t), a.h.ptr, tab, length(tab)))
end
"`g + ∂ₜuBC` on every index at t₀ (`which`=0) or t₁ (`which`=1) for accelerate! (src/Flow.jl:69-73)."
function set_accel_field!(a::B200Flow{D,T}, which::Integer, f::Function, t) where {D,T}
    Ng = size(a.p); fs = slab_fn(f, a.koff, Val(D), T)
    fld = T[fs(i, WaterLily.loc(i, I, T), T(t)) for I in CartesianIndices(Ng), i in 1:D]
    check(a.h, ccall((:wl_set_accel_field, LIB), Cint, (Ptr{Cvoid}, Cint, Ptr{Cfloat}, Csize_t), a.h.ptr, which, fld, length(fld)))
end

# ---- MeanFlow (src/Metrics.jl:205-263): device-resident running averages --------------------------------
struct B200MeanFlow{T}
    h::Handle; P::B200Array{T}; U::B200Array{T}; UU::Union{Nothing,B200Array{T}}; t::Vector{T}; uu_stats::Bool
end
function B200MeanFlow(flow::B200Flow{D,T}; t_init=time(flow), uu_stats=false) where {D,T}
    check(flow.h, ccall((:wl_meanflow_init, LIB), Cint, (Ptr{Cvoid}, Cint), flow.h.ptr, uu_stats))
    ng = size(flow.p); a(f, d) = B200Array{T,length(d)}(flow.h, Int32(f), Int32(0), d)
    B200MeanFlow{T}(flow.h, a(20, ng), a(21, (ng..., D)), uu_stats ? a(22, (ng..., D, D)) : nothing, T[t_init], uu_stats)
end
time(m::B200MeanFlow) = m.t[end] - m.t[1]
function update!(m::B200MeanFlow{T}, flow::B200Flow) where T
    dt = time(flow) - m.t[end]
    ε = length(m.t) == 1 ? one(T) : dt / (dt + time(m) + eps(T))
    check(m.h, ccall((:wl_meanflow_update, LIB), Cint, (Ptr{Cvoid}, Cfloat), m.h.ptr, ε))
    push!(m.t, m.t[end] + dt)
end
function WaterLily.reset!(m::B200MeanFlow{T}; t_init=0.0) where T      # WaterLily's own generic (src/Metrics.jl:229), not exported
    check(m.h, ccall((:wl_meanflow_init, LIB), Cint, (Ptr{Cvoid}, Cint), m.h.ptr, m.uu_stats))
    empty!(m.t); push!(m.t, T(t_init))
end
WaterLily.copy!(flow::B200Flow, m::B200MeanFlow) = check(m.h, ccall((:wl_meanflow_copy_to_flow, LIB), Cint, (Ptr{Cvoid},), m.h.ptr))

"Drop-in constructor: WaterLily's own `Simulation` with the two factories swapped."
B200Simulation(dims, uBC, L; device=0, kw...) =
    WaterLily.Simulation(dims, uBC, L; kw..., flow_ctor=(d, u; k...) -> B200Flow(d, u; device, k...), pois_ctor=B200Poisson)

export B200Simulation, B200Flow, B200Poisson, B200Array, B200Sphere, B200Cylinder, B200Torus
end # module
