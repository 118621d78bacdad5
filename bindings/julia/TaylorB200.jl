# TaylorB200.jl -- ccall shim over libtaylorb200.so (include/taylorb200.h).
# UNTESTED in this repository's CI: the build image has no Julia. It mirrors, entry point for entry point,
# the tested Python ctypes host (taylorintegration.jl_b200/_lib.py, api.py, tape.py).
#
# Replaces, for ensembles, TaylorIntegration.taylorinteg (src/integrator/taylorinteg.jl:306-381):
#     taylorinteg(tape, q0::Matrix (N x dof), t0, tmax, order, abstol, params; maxsteps=500, ...)
# The tape is produced by tracing the RHS with `Ser` values (what @taylorize's parser would emit,
# src/parse_eqs.jl:963-1135), e.g.
#     tape = taylorize(4, 1) do dq, q, p, t
#         r = (q[1]^2 + q[2]^2)^1.5
#         dq[1] = q[3]; dq[2] = q[4]; dq[3] = p[1]*q[1]/r; dq[4] = p[1]*q[2]/r
#     end
module TaylorB200

export taylorize, taylorinteg, taylorinteg_events, taylorinteg_trange, free_kernels, Tape

const LIB = get(ENV, "TAYLORB200_LIB", "libtaylorb200.so")

# ---- ABI mirrors (field order = include/taylorb200.h) ----------------------------------------------------
struct TbOp
    op::Int32; res::Int32; a::Int32; b::Int32; aux::Int32; flags::Int32; imm::Float64
end
struct TbJacEntry
    kind::Int32; operand::Int32; c::Float64
end
struct TbOdeDesc
    dof::Int32; npar::Int32; ntmp::Int32; nops::Int32; ev_nops::Int32; jac_nops::Int32
    ops::Ptr{TbOp}
    ev_g::Int32; ev_gate::Int32; ev_gate_cmp::Int32; has_jac::Int32
    jac::Ptr{TbJacEntry}
end
mutable struct TbCompileOpts
    order::Int32; mode::Int32; kind::Int32; block_threads::Int32
    reg_budget::Int32; min_blocks_per_sm::Int32; flags::Int32; tmem_budget::Int32
    TbCompileOpts() = new(0, 0, 0, 0, 0, 0, 0, 0)
end
mutable struct TbOpts
    abstol::Float64; reltol::Float64; minstepsize::Float64; maxstepsize::Float64
    maxsteps::Int64
    eventorder::Int32; newtoniter::Int32
    nrabstol::Float64
    event_capacity::Int32; params_per_traj::Int32
    steps_stride::Int64
    sort_by_cost::Int32; reserved::Int32
    TbOpts() = new(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0)
end
mutable struct TbOut
    xfinal::Ptr{Float64}; tfinal::Ptr{Float64}; nsteps::Ptr{Int64}; status::Ptr{Int32}
    xv_trange::Ptr{Float64}; tv::Ptr{Float64}; xv::Ptr{Float64}; psol::Ptr{Float64}
    tevents::Ptr{Float64}; xevents::Ptr{Float64}; gresids::Ptr{Float64}; nevents::Ptr{Int32}
    lambda::Ptr{Float64}; lambda_v::Ptr{Float64}
    lyap_state::Ptr{Float64}; lyap_state0::Ptr{Float64}   # Lyapunov continue-from-state (out / in)
    TbOut() = new(ntuple(_ -> C_NULL, 16)...)
end

const OP = (IDENTITY = 1, NEG = 2, ADD = 3, SUB = 4, MUL = 5, DIV = 6, SQR = 7, SQRT = 8, POW = 9, EXP = 10,
            LOG = 11, SINCOS = 12, TAN = 15, SINHCOSH = 16, TANH = 17, ATAN = 18, ASIN = 19, ACOS = 20)
const K_NONE, K_X, K_DX, K_TMP, K_PAR, K_TIME, K_IMM = 0, 1, 2, 3, 4, 5, 6
operand(kind, index) = Int32((UInt32(kind) << 24) | (UInt32(index) & 0x00ffffff))
const ST_MAXSTEPS, ST_ZERO_STEP, ST_NR_NOCONV = 1, 2, 4

lasterror() = unsafe_string(ccall((:tb_last_error, LIB), Cstring, ()))

# ---- tracer: the tape @taylorize would emit ------------------------------------------------------------
mutable struct Tape
    dof::Int; npar::Int; ntmp::Int; ops::Vector{TbOp}
end
struct Ser           # a series-valued operand while tracing
    tape::Tape; ref::Int32
end
struct Par           # a scalar parameter (`local μ = p[1]`)
    tape::Tape; ref::Int32
end
newtmp(t::Tape) = (t.ntmp += 1; Ser(t, operand(K_TMP, t.ntmp - 1)))
ref(x::Ser) = (x.ref, 0.0); ref(x::Par) = (x.ref, 0.0); ref(x::Real) = (operand(K_IMM, 0), Float64(x))
function emit(t::Tape, op, a, b = nothing; aux = Int32(0), imm = nothing)
    r = newtmp(t)
    (ra, ia) = ref(a); (rb, ib) = b === nothing ? (Int32(0), 0.0) : ref(b)
    v = imm !== nothing ? Float64(imm) : (a isa Real ? ia : ib)
    push!(t.ops, TbOp(op, r.ref, ra, rb, aux, 0, v)); r
end
const Operand = Union{Ser, Par, Real}
for (f, op) in ((:+, OP.ADD), (:-, OP.SUB), (:*, OP.MUL), (:/, OP.DIV))
    @eval Base.$f(a::Ser, b::Operand) = emit(a.tape, $op, a, b)
    @eval Base.$f(a::Union{Par, Real}, b::Ser) = emit(b.tape, $op, a, b)
end
Base.:-(a::Ser) = emit(a.tape, OP.NEG, a)
Base.:^(a::Ser, n::Real) = emit(a.tape, OP.POW, a; aux = newtmp(a.tape).ref, imm = n)
Base.:^(a::Ser, n::Integer) = emit(a.tape, OP.POW, a; aux = newtmp(a.tape).ref, imm = n)
Base.sqrt(a::Ser) = emit(a.tape, OP.SQRT, a)
Base.exp(a::Ser) = emit(a.tape, OP.EXP, a)
Base.log(a::Ser) = emit(a.tape, OP.LOG, a)
function Base.sin(a::Ser); c = newtmp(a.tape); emit(a.tape, OP.SINCOS, a; aux = c.ref); end
function Base.cos(a::Ser)
    t = a.tape; s = newtmp(t); c = newtmp(t)
    push!(t.ops, TbOp(OP.SINCOS, s.ref, a.ref, 0, c.ref, 0, 0.0)); c
end
struct DxSink; tape::Tape; end
function Base.setindex!(d::DxSink, v, i::Integer)   # dq[i] = <expr>  ->  identity!(dq[i], expr) or retarget
    t = d.tape; (rv, iv) = ref(v)
    push!(t.ops, TbOp(OP.IDENTITY, operand(K_DX, i - 1), rv, 0, 0, 0, iv))
end
"taylorize(f!, dof, npar): trace f!(dq, q, p, t) into a Tape"
function taylorize(f!, dof::Integer, npar::Integer)
    t = Tape(dof, npar, 0, TbOp[])
    q = [Ser(t, operand(K_X, i - 1)) for i in 1:dof]
    p = [Par(t, operand(K_PAR, j - 1)) for j in 1:npar]
    f!(DxSink(t), q, p, Ser(t, operand(K_TIME, 0)))
    t
end

# ---- compile (cached) --------------------------------------------------------------------------------
# One kernel per (tape, order, mode, kind, flags); the ode handle is freed as soon as tb_compile has copied it.
const KIND_PLAIN, KIND_EVENTS = 0, 1
const CF_FINAL_ONLY = 262144
const KERNELS = Dict{Tuple{UInt,Int,Int,Int,Int,Int32},Ptr{Cvoid}}()
function compile(t::Tape, order::Integer; mode = 0, kind = 0, flags = 0, ev_g::Int32 = Int32(0))
    key = (objectid(t), Int(order), Int(mode), Int(kind), Int(flags), ev_g)
    haskey(KERNELS, key) && return KERNELS[key]
    ode = GC.@preserve t begin
        desc = TbOdeDesc(t.dof, t.npar, t.ntmp, length(t.ops), 0, 0, pointer(t.ops), ev_g, 0, 0, 0, C_NULL)
        ccall((:tb_ode_create, LIB), Ptr{Cvoid}, (Ref{TbOdeDesc},), desc)
    end
    ode == C_NULL && error("tb_ode_create: " * lasterror())
    co = TbCompileOpts()
    ccall((:tb_compile_opts_default, LIB), Cvoid, (Ref{TbCompileOpts},), co)
    co.order, co.mode, co.kind, co.flags = order, mode, kind, flags
    k = ccall((:tb_compile, LIB), Ptr{Cvoid}, (Ptr{Cvoid}, Ref{TbCompileOpts}), ode, co)
    err = k == C_NULL ? lasterror() : ""
    ccall((:tb_ode_free, LIB), Cvoid, (Ptr{Cvoid},), ode)
    k == C_NULL && error("tb_compile: " * err)
    KERNELS[key] = k
end
"free every cached kernel (tb_kernel_free)"
function free_kernels()
    for k in values(KERNELS); ccall((:tb_kernel_free, LIB), Cvoid, (Ptr{Cvoid},), k); end
    empty!(KERNELS)
end

function make_opts(abstol, reltol, minstepsize, maxstepsize, maxsteps, params)
    o = TbOpts(); ccall((:tb_opts_default, LIB), Cvoid, (Ref{TbOpts},), o)
    o.abstol, o.reltol, o.minstepsize, o.maxstepsize, o.maxsteps = abstol, reltol, minstepsize, maxstepsize, maxsteps
    o.params_per_traj = params isa Matrix ? 1 : 0
    o
end
function warn_status(st)  # the reference's warnings: taylorinteg.jl:360-363, 372-377; rootfinding.jl:222-224
    any(s -> s & ST_MAXSTEPS != 0, st) && @warn "Maximum number of integration steps reached; exiting."
    any(s -> s & ST_ZERO_STEP != 0, st) && @warn "The step-size is zero; aborting integration."
    any(s -> s & ST_NR_NOCONV != 0, st) && @warn "Newton-Raphson did not converge for prescribed tolerance and maximum allowed iterations."
end

"""
    taylorinteg(tape, q0::Matrix{Float64}, t0, tmax, order, abstol, params=Float64[]; maxsteps=500, reltol=0,
                minstepsize=0, maxstepsize=Inf, fast=false)

Ensemble counterpart of `TaylorIntegration.taylorinteg` (taylorinteg.jl:306-327). `q0` is `N x dof` (column
major = the ABI's [dof x N] layout). `params` is `npar` (shared) or `N x npar` (swept). Returns
`(xfinal::Matrix N x dof, tfinal, nsteps, status)`.
"""
function taylorinteg(tape::Tape, q0::Matrix{Float64}, t0::Real, tmax::Real, order::Integer, abstol::Real,
                     params::VecOrMat{Float64} = Float64[]; maxsteps::Integer = 500, reltol::Real = 0.0,
                     minstepsize::Real = 0.0, maxstepsize::Real = Inf, fast::Bool = false,
                     devices = nothing, block::Integer = 4096)
    N, dof = size(q0); dof == tape.dof || throw(ArgumentError("q0 must be N x dof"))
    k = compile(tape, order; mode = fast ? 1 : 0, flags = CF_FINAL_ONLY)
    o = make_opts(abstol, reltol, minstepsize, maxstepsize, maxsteps, params)
    xf = Matrix{Float64}(undef, N, dof); tf = Vector{Float64}(undef, N)
    ns = Vector{Int64}(undef, N); st = Vector{Int32}(undef, N)
    out = TbOut(); out.xfinal, out.tfinal, out.nsteps, out.status = pointer(xf), pointer(tf), pointer(ns), pointer(st)
    rc = GC.@preserve xf tf ns st q0 params begin
        if devices === nothing
            ccall((:tb_integrate, LIB), Cint,
                  (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Float64, Float64, Ref{TbOpts}, Ref{TbOut}),
                  k, N, q0, params, t0, tmax, o, out)
        else  # one ensemble over several GPUs of this process (tb_integrate_multi, round-robin blocks)
            dv = Int32.(devices)
            ccall((:tb_integrate_multi, LIB), Cint,
                  (Ptr{Cvoid}, Ptr{Int32}, Int32, Int64, Int64, Ptr{Float64}, Ptr{Float64}, Float64, Float64,
                   Ptr{Float64}, Int64, Ref{TbOpts}, Ref{TbOut}),
                  k, dv, length(dv), block, N, q0, params, t0, tmax, C_NULL, 0, o, out)
        end
    end
    rc == 0 || error("tb_integrate: " * lasterror())
    warn_status(st)
    (xf, tf, ns, st)
end

"""
    taylorinteg_trange(tape, q0, trange, order, abstol, params=Float64[]; maxsteps=500, ...)

`taylorinteg(f!, q0, trange, order, abstol, params)` (src/integrator/taylorinteg.jl:558-642) over an ensemble:
returns `(xv::Array{Float64,3} (N x nt x dof, NaN past the last time reached), xfinal, nsteps, status)`.
"""
function taylorinteg_trange(tape::Tape, q0::Matrix{Float64}, trange::AbstractVector{<:Real}, order::Integer,
                            abstol::Real, params::VecOrMat{Float64} = Float64[]; maxsteps::Integer = 500,
                            reltol::Real = 0.0, minstepsize::Real = 0.0, maxstepsize::Real = Inf, fast::Bool = false)
    N, dof = size(q0); dof == tape.dof || throw(ArgumentError("q0 must be N x dof"))
    tr = collect(Float64, trange); nt = length(tr)
    @assert nt >= 2 && (issorted(tr) || issorted(tr, rev = true))     # taylorinteg.jl:489,573
    k = compile(tape, order; mode = fast ? 1 : 0)
    o = make_opts(abstol, reltol, minstepsize, maxstepsize, maxsteps, params)
    xv = Array{Float64,3}(undef, N, nt, dof); xf = Matrix{Float64}(undef, N, dof)
    ns = Vector{Int64}(undef, N); st = Vector{Int32}(undef, N)
    out = TbOut(); out.xv_trange, out.xfinal, out.nsteps, out.status = pointer(xv), pointer(xf), pointer(ns), pointer(st)
    rc = GC.@preserve xv xf ns st q0 params tr ccall((:tb_integrate_trange, LIB), Cint,
        (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Ptr{Float64}, Int64, Ref{TbOpts}, Ref{TbOut}),
        k, N, q0, params, tr, nt, o, out)
    rc == 0 || error("tb_integrate_trange: " * lasterror())
    warn_status(st)
    (xv, xf, ns, st)
end

"""
    taylorinteg_events(tape, gcomp, q0, t0, tmax, order, abstol, params=Float64[]; eventorder=0, newtoniter=10,
                       nrabstol=eps(), event_capacity=maxsteps+1, ...)

`taylorinteg(f!, g, q0, t0, tmax, order, abstol, params)` (src/rootfinding.jl:301-438) for an event function that is a
state component, `g(dx, x, p, t) = (true, x[gcomp])` (test/rootfinding.jl:26).  Returns
`(xfinal, tfinal, nsteps, status, tevents (N x cap), xevents (N x cap x dof), gresids (N x cap), nevents)`; only the
first `nevents[i]` entries of row i are meaningful.
"""
function taylorinteg_events(tape::Tape, gcomp::Integer, q0::Matrix{Float64}, t0::Real, tmax::Real, order::Integer,
                            abstol::Real, params::VecOrMat{Float64} = Float64[]; maxsteps::Integer = 500,
                            eventorder::Integer = 0, newtoniter::Integer = 10, nrabstol::Real = eps(Float64),
                            event_capacity::Integer = maxsteps + 1, reltol::Real = 0.0, minstepsize::Real = 0.0,
                            maxstepsize::Real = Inf, fast::Bool = false)
    @assert order >= eventorder                                        # rootfinding.jl:321
    N, dof = size(q0); dof == tape.dof || throw(ArgumentError("q0 must be N x dof"))
    k = compile(tape, order; mode = fast ? 1 : 0, kind = KIND_EVENTS, ev_g = operand(K_X, gcomp - 1))
    o = make_opts(abstol, reltol, minstepsize, maxstepsize, maxsteps, params)
    o.eventorder, o.newtoniter, o.nrabstol, o.event_capacity = eventorder, newtoniter, nrabstol, event_capacity
    cap = Int(event_capacity)
    xf = Matrix{Float64}(undef, N, dof); tf = Vector{Float64}(undef, N)
    ns = Vector{Int64}(undef, N); st = Vector{Int32}(undef, N); ne = Vector{Int32}(undef, N)
    te = zeros(N, cap); xe = zeros(N, cap, dof); gr = zeros(N, cap)
    out = TbOut()
    out.xfinal, out.tfinal, out.nsteps, out.status = pointer(xf), pointer(tf), pointer(ns), pointer(st)
    out.tevents, out.xevents, out.gresids, out.nevents = pointer(te), pointer(xe), pointer(gr), pointer(ne)
    rc = GC.@preserve xf tf ns st ne te xe gr q0 params ccall((:tb_integrate, LIB), Cint,
        (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Float64, Float64, Ref{TbOpts}, Ref{TbOut}),
        k, N, q0, params, t0, tmax, o, out)
    rc == 0 || error("tb_integrate: " * lasterror())
    warn_status(st)
    (xf, tf, ns, st, te, xe, gr, ne)
end

# lyap_taylorinteg (src/lyapunovspectrum.jl:293-338): the column-parallel kernel (TB_KIND_LYAPCOL) integrates the column
# system [x; v], v' = J(x) v, which the Python host derives from the tape by first-order forward mode
# (taylorintegration.jl_b200/tape.py lyap_column_tape).  A Julia caller passes that tape (same TbOp records) to
# `compile(tape, order; kind = 3)` and calls tb_integrate with `out.lambda`; the tracer above does not build it yet.

end # module
