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

export taylorize, taylorinteg, Tape

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
    reg_budget::Int32; min_blocks_per_sm::Int32; flags::Int32; reserved::Int32
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
    TbOut() = new(ntuple(_ -> C_NULL, 14)...)
end

const OP = (IDENTITY = 1, NEG = 2, ADD = 3, SUB = 4, MUL = 5, DIV = 6, SQR = 7, SQRT = 8, POW = 9, EXP = 10,
            LOG = 11, SINCOS = 12)
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

# ---- compile + integrate ---------------------------------------------------------------------------------
function compile(t::Tape, order::Integer; mode = 0, kind = 0)
    GC.@preserve t begin
        desc = TbOdeDesc(t.dof, t.npar, t.ntmp, length(t.ops), 0, 0, pointer(t.ops), 0, 0, 0, 0, C_NULL)
        ode = ccall((:tb_ode_create, LIB), Ptr{Cvoid}, (Ref{TbOdeDesc},), desc)
    end
    ode == C_NULL && error("tb_ode_create: " * lasterror())
    co = TbCompileOpts()
    ccall((:tb_compile_opts_default, LIB), Cvoid, (Ref{TbCompileOpts},), co)
    co.order, co.mode, co.kind = order, mode, kind
    k = ccall((:tb_compile, LIB), Ptr{Cvoid}, (Ptr{Cvoid}, Ref{TbCompileOpts}), ode, co)
    k == C_NULL && error("tb_compile: " * lasterror())
    k
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
                     minstepsize::Real = 0.0, maxstepsize::Real = Inf, fast::Bool = false)
    N, dof = size(q0); dof == tape.dof || throw(ArgumentError("q0 must be N x dof"))
    k = compile(tape, order; mode = fast ? 1 : 0)
    o = TbOpts(); ccall((:tb_opts_default, LIB), Cvoid, (Ref{TbOpts},), o)
    o.abstol, o.reltol, o.minstepsize, o.maxstepsize, o.maxsteps = abstol, reltol, minstepsize, maxstepsize, maxsteps
    o.params_per_traj = params isa Matrix ? 1 : 0
    xf = Matrix{Float64}(undef, N, dof); tf = Vector{Float64}(undef, N)
    ns = Vector{Int64}(undef, N); st = Vector{Int32}(undef, N)
    out = TbOut(); out.xfinal, out.tfinal, out.nsteps, out.status = pointer(xf), pointer(tf), pointer(ns), pointer(st)
    rc = GC.@preserve xf tf ns st ccall((:tb_integrate, LIB), Cint,
        (Ptr{Cvoid}, Int64, Ptr{Float64}, Ptr{Float64}, Float64, Float64, Ref{TbOpts}, Ref{TbOut}),
        k, N, q0, params, t0, tmax, o, out)
    ccall((:tb_kernel_free, LIB), Cvoid, (Ptr{Cvoid},), k)
    rc == 0 || error("tb_integrate: " * lasterror())
    any(s -> s & ST_MAXSTEPS != 0, st) && @warn "Maximum number of integration steps reached; exiting." # taylorinteg.jl:372-377
    any(s -> s & ST_ZERO_STEP != 0, st) && @warn "The step-size is zero; aborting integration."          # taylorinteg.jl:360-363
    (xf, tf, ns, st)
end

end # module
