# TCIB200.jl — reference-side binding of libtci_b200.so (include/tci_b200.h).
#
# This is the shim a maintainer of TensorCrossInterpolation.jl adds to route the TensorCI2 two-site-sweep hot path
# through the B200 library.  It overloads the reference's own extension points and nothing else:
#   B1  (f::BatchEvaluator{Float64})(leftindexset, rightindexset, Val(M))   src/batcheval.jl:196-214, docs/src/index.md:181-237
#   B2  rrlu / rrlu! / MatrixLUCI on a device-resident Pi matrix            src/matrixlu.jl:420-463, src/matrixluci.jl:55
#   B3  left/right(::MatrixLUCI), the solve of setsitetensor!, _contract,   src/matrixluci.jl:256-283, src/tensorci2.jl:599-629,
#       sum, evaluate                                                        src/contraction.jl:193-215, src/abstracttensortrain.jl:328-441
#   and, optionally, the bond loop itself: halfsweep! / fillsitetensors! as single launches (tci_dsweep_*, csrc/sweep.cu)
# Sweep strategy, index-set bookkeeping and convergence control stay in the unmodified Julia package.
#
# NOT TESTED in this repository: the build image and the GPU box have no Julia (see DESIGN.md §1).  The same C ABI is
# exercised from Python (tensorcrossinterpolation.jl_b200/_lib.py), which mirrors this file call for call.
module TCIB200

import TensorCrossInterpolation as TCI
import TensorCrossInterpolation: MultiIndex, BatchEvaluator, rrLU

const LIB = get(ENV, "TCI_B200_LIB", "libtci_b200.so")

# ---- status codes -> the reference's exceptions (include/tci_b200.h: tci_status) ---------------------------------------
const TCI_OK = Cint(0)
function check(h::Ptr{Cvoid}, rc::Cint)
    rc == TCI_OK && return nothing
    msg = unsafe_string(ccall((:tci_last_error, LIB), Cstring, (Ptr{Cvoid},), h))
    rc == -1 && throw(ArgumentError(msg))          # TCI_ERR_INVALID
    error(msg)                                     # "lu.L contains NaNs", "Pivot matrix at bond b is not square!", CUDA errors ...
end

# ---- handle: one CUDA device + stream + workspaces (one per Julia task / per GPU) ---------------------------------------
mutable struct Handle
    ptr::Ptr{Cvoid}
    function Handle(device::Integer=0)
        r = Ref{Ptr{Cvoid}}(C_NULL)
        rc = ccall((:tci_create, LIB), Cint, (Ref{Ptr{Cvoid}}, Cint), r, device)
        rc == TCI_OK || error(unsafe_string(ccall((:tci_last_error, LIB), Cstring, (Ptr{Cvoid},), C_NULL)))
        h = new(r[])
        finalizer(x -> ccall((:tci_destroy, LIB), Cint, (Ptr{Cvoid},), x.ptr), h)
        return h
    end
end
const DEFAULT = Ref{Union{Nothing,Handle}}(nothing)
handle() = (DEFAULT[] === nothing && (DEFAULT[] = Handle(0)); DEFAULT[]::Handle)

# ---- device matrix (column-major FP64, leading dimension padded to 128 B) ------------------------------------------------
mutable struct DeviceMatrix <: AbstractMatrix{Float64}
    h::Handle
    ptr::Ptr{Float64}
    m::Int
    n::Int
    ld::Int
    function DeviceMatrix(h::Handle, m::Int, n::Int)
        ld = max(cld(m, 16) * 16, 1)
        p = Ref{Ptr{Cvoid}}(C_NULL)
        check(h.ptr, ccall((:tci_malloc, LIB), Cint, (Ptr{Cvoid}, Ref{Ptr{Cvoid}}, Csize_t), h.ptr, p, 8 * ld * max(n, 1)))
        A = new(h, Ptr{Float64}(p[]), m, n, ld)
        finalizer(x -> ccall((:tci_free, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}), x.h.ptr, x.ptr), A)
        return A
    end
end
Base.size(A::DeviceMatrix) = (A.m, A.n)
function Base.Array(A::DeviceMatrix)
    out = Matrix{Float64}(undef, A.ld, A.n)
    check(A.h.ptr, ccall((:tci_memcpy_d2h, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Csize_t), A.h.ptr, out, A.ptr, 8 * A.ld * A.n))
    return out[1:A.m, :]
end
Base.getindex(A::DeviceMatrix, i::Int, j::Int) = Array(A)[i, j]      # debugging only; the sweep never indexes Pi element-wise

# ---- B1: device functor as a BatchEvaluator ------------------------------------------------------------------------------
"""
    B200Function(localdims; kind=:lorentz, params=[1.0])        built-in functor (include/tci_b200.h: TCI_F_*_ID)
    B200Function(localdims, cuda_source; params=Float64[])      plugin: CUDA source defining
        __device__ double tci_user_f(const int* idx, int n, const double* params, int nparams);
"""
mutable struct B200Function <: BatchEvaluator{Float64}
    h::Handle
    f::Ptr{Cvoid}
    localdims::Vector{Int32}
end
const KINDS = Dict(:lorentz => 1, :qexpsin => 2, :qexpplussin => 3, :sum => 4, :qexp2 => 5)
function B200Function(localdims::Vector{Int}; kind::Symbol=:lorentz, params::Vector{Float64}=[1.0], h::Handle=handle())
    r = Ref{Ptr{Cvoid}}(C_NULL)
    check(h.ptr, ccall((:tci_functor_builtin, LIB), Cint, (Ptr{Cvoid}, Cint, Ptr{Float64}, Cint, Ref{Ptr{Cvoid}}),
                       h.ptr, KINDS[kind], params, length(params), r))
    return _finalize(B200Function(h, r[], Int32.(localdims)))
end
function B200Function(localdims::Vector{Int}, cuda_source::String; params::Vector{Float64}=Float64[], h::Handle=handle())
    r = Ref{Ptr{Cvoid}}(C_NULL)
    check(h.ptr, ccall((:tci_functor_from_source, LIB), Cint, (Ptr{Cvoid}, Cstring, Ptr{Float64}, Cint, Ref{Ptr{Cvoid}}),
                       h.ptr, cuda_source, params, length(params), r))
    return _finalize(B200Function(h, r[], Int32.(localdims)))
end
_finalize(f::B200Function) = (finalizer(x -> ccall((:tci_functor_destroy, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}), x.h.ptr, x.f), f); f)

_table(set::Vector{MultiIndex}) = isempty(set) ? Matrix{Int32}(undef, 0, 0) : Int32[set[i][p] for p in 1:length(first(set)), i in 1:length(set)]  # (len, count) column-major == count rows of len ints

# batch evaluation into a device matrix (what updatepivots! wants: Pi never visits the host)
function evaluate_device!(f::B200Function, out::DeviceMatrix, left::Vector{MultiIndex}, right::Vector{MultiIndex},
                          stride_c::Int, stride_j::Int, maxabs::Ref{Float64})
    L, R = _table(left), _table(right)
    check(f.h.ptr, ccall((:tci_batcheval_f64, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Int32}, Cint, Ptr{Int32}, Int64, Cint, Ptr{Int32}, Int64, Cint, Ptr{Float64}, Int64, Int64, Ref{Float64}),
        f.h.ptr, f.f, f.localdims, length(f.localdims), L, length(left), size(L, 1), R, length(right), size(R, 1),
        out.ptr, stride_c, stride_j, maxabs))
    return out
end

# the reference's protocol: returns a fresh host Array{Float64,M+2} (src/batcheval.jl:196-214)
function (f::B200Function)(left::Vector{MultiIndex}, right::Vector{MultiIndex}, ::Val{M})::Array{Float64,M + 2} where {M}
    if length(left) * length(right) == 0
        return Array{Float64,M + 2}(undef, ntuple(i -> 0, M + 2)...)
    end
    nl = length(first(left))
    cdims = Int.(f.localdims[nl+1:nl+M])
    rows = length(left) * prod(cdims; init=1)
    tmp = DeviceMatrix(f.h, rows, length(right))
    evaluate_device!(f, tmp, left, right, length(left), tmp.ld, Ref(0.0))
    return reshape(Array(tmp), length(left), cdims..., length(right))
end
(f::B200Function)(indexset::MultiIndex)::Float64 = f([indexset[1:1]], [indexset[2:end]], Val(0))[1, 1]

# ---- B2: rrlu on host matrices (drop-in for TCI.rrlu) and on device matrices (used by the sweep) ---------------------------
function rrlu_b200(A::Matrix{Float64}; maxrank::Int=typemax(Int), reltol::Number=1e-14, abstol::Number=0.0,
                   leftorthogonal::Bool=true, h::Handle=handle())::rrLU{Float64}
    m, n = size(A)
    mr = min(maxrank, m, n)
    rowperm, colperm = Vector{Int}(undef, m), Vector{Int}(undef, n)
    L, U = Matrix{Float64}(undef, m, max(mr, 1)), Matrix{Float64}(undef, max(mr, 1), n)
    npivot, err = Ref{Int64}(0), Ref{Float64}(0.0)
    check(h.ptr, ccall((:tci_rrlu_host_f64, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Int64, Int64, Int64, Int64, Float64, Float64, Cint, Ptr{Int64}, Ptr{Int64}, Ref{Int64}, Ref{Float64},
         Ptr{Float64}, Int64, Ptr{Float64}, Int64, Ptr{Float64}),
        h.ptr, A, m, n, max(m, 1), mr, reltol, abstol, leftorthogonal, rowperm, colperm, npivot, err, L, max(m, 1), U, max(mr, 1), C_NULL))
    r = npivot[]
    return rrLU{Float64}(rowperm, colperm, L[:, 1:r], U[1:r, :], leftorthogonal, r, err[])
end

"""Device-side MatrixLUCI: rrlu! in place on Pi, factors formed on the device (tci_luci_factors_f64)."""
struct DeviceLUCI
    Pi::DeviceMatrix
    rowperm::Vector{Int}
    colperm::Vector{Int}
    L::DeviceMatrix
    U::DeviceMatrix
    leftorthogonal::Bool
    npivot::Int
    error::Float64
    pivoterrors::Vector{Float64}
end
function DeviceLUCI(Pi::DeviceMatrix; maxrank::Int=typemax(Int), reltol::Number=1e-14, abstol::Number=0.0, leftorthogonal::Bool=true)
    h, m, n = Pi.h, Pi.m, Pi.n
    mr = min(maxrank, m, n)
    rowperm, colperm = Vector{Int}(undef, m), Vector{Int}(undef, n)
    L, U = DeviceMatrix(h, m, max(mr, 1)), DeviceMatrix(h, max(mr, 1), n)
    npivot, err, pe = Ref{Int64}(0), Ref{Float64}(0.0), Vector{Float64}(undef, mr + 1)
    check(h.ptr, ccall((:tci_rrlu_inplace_f64, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Int64, Int64, Int64, Int64, Float64, Float64, Cint, Ptr{Int64}, Ptr{Int64}, Ref{Int64}, Ref{Float64},
         Ptr{Float64}, Int64, Ptr{Float64}, Int64, Ptr{Float64}),
        h.ptr, Pi.ptr, m, n, Pi.ld, mr, reltol, abstol, leftorthogonal, rowperm, colperm, npivot, err, L.ptr, L.ld, U.ptr, U.ld, pe))
    return DeviceLUCI(Pi, rowperm, colperm, L, U, leftorthogonal, npivot[], err[], pe[1:npivot[]+1])
end
TCI.rowindices(l::DeviceLUCI) = l.rowperm[1:l.npivot]
TCI.colindices(l::DeviceLUCI) = l.colperm[1:l.npivot]
TCI.npivots(l::DeviceLUCI) = l.npivot
TCI.pivoterrors(l::DeviceLUCI) = l.pivoterrors
function factors(l::DeviceLUCI)      # left(luci), right(luci)  src/matrixluci.jl:256-283
    h, m, n, r = l.Pi.h, l.Pi.m, l.Pi.n, l.npivot
    left, right = DeviceMatrix(h, m, r), DeviceMatrix(h, r, n)
    check(h.ptr, ccall((:tci_luci_factors_f64, LIB), Cint,
        (Ptr{Cvoid}, Int64, Int64, Int64, Cint, Ptr{Int64}, Ptr{Int64}, Ptr{Float64}, Int64, Ptr{Float64}, Int64, Ptr{Float64}, Int64, Ptr{Float64}, Int64),
        h.ptr, m, n, r, l.leftorthogonal, l.rowperm, l.colperm, l.L.ptr, l.L.ld, l.U.ptr, l.U.ld, left.ptr, left.ld, right.ptr, right.ld))
    return left, right
end
TCI.left(l::DeviceLUCI) = Array(factors(l)[1])
TCI.right(l::DeviceLUCI) = Array(factors(l)[2])

# ---- the one method of the sweep that is re-routed: the :full branch of updatepivots! (src/tensorci2.jl:845-868) ----------
# Identical to the reference except that Pi is filled and factorised on the device.
function TCI.updatepivots!(tci::TCI.TensorCI2{Float64}, b::Int, f::B200Function, leftorthogonal::Bool;
                           reltol::Float64=1e-14, abstol::Float64=0.0, maxbonddim::Int=typemax(Int),
                           sweepdirection::Symbol=:forward, pivotsearch::Symbol=:full, verbosity::Int=0,
                           extraIset::Vector{MultiIndex}=MultiIndex[], extraJset::Vector{MultiIndex}=MultiIndex[])
    pivotsearch === :full || error("B200 path implements pivotsearch=:full")
    TCI.invalidatesitetensors!(tci)
    Icombined = union(TCI.kronecker(tci.Iset[b], tci.localdims[b]), extraIset)
    Jcombined = union(TCI.kronecker(tci.localdims[b+1], tci.Jset[b+1]), extraJset)
    Pi = DeviceMatrix(f.h, length(Icombined), length(Jcombined))
    mx = Ref(tci.maxsamplevalue)
    evaluate_device!(f, Pi, Icombined, Jcombined, length(Icombined), Pi.ld, mx)     # filltensor + updatemaxsample!
    tci.maxsamplevalue = mx[]
    luci = DeviceLUCI(Pi; reltol=reltol, abstol=abstol, maxrank=maxbonddim, leftorthogonal=leftorthogonal)
    tci.Iset[b+1] = Icombined[TCI.rowindices(luci)]
    tci.Jset[b] = Jcombined[TCI.colindices(luci)]
    if length(extraIset) == 0 && length(extraJset) == 0
        TCI.setsitetensor!(tci, b, TCI.left(luci))
        TCI.setsitetensor!(tci, b + 1, TCI.right(luci))
    end
    TCI.updateerrors!(tci, b, TCI.pivoterrors(luci))
    nothing
end

# ---- B3: TensorTrain sum / evaluate on device-resident cores ---------------------------------------------------------------
struct DeviceTT
    h::Handle
    cores::Vector{DeviceMatrix}          # (chi_l*d) x chi_r each
    shapes::Vector{Int64}                # 3 per core
end
function DeviceTT(tt::TCI.AbstractTensorTrain{Float64}; h::Handle=handle())
    cores, shapes = DeviceMatrix[], Int64[]
    for T in TCI.sitetensors(tt)
        chil, chir = size(T, 1), size(T)[end]
        d = div(length(T), chil * chir)
        M = DeviceMatrix(h, chil * d, chir)
        host = zeros(M.ld, chir); host[1:chil*d, :] = reshape(T, chil * d, chir)
        check(h.ptr, ccall((:tci_memcpy_h2d, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Csize_t), h.ptr, M.ptr, host, 8 * length(host)))
        push!(cores, M); append!(shapes, (chil, d, chir))
    end
    return DeviceTT(h, cores, shapes)
end
function Base.sum(tt::DeviceTT)
    out = Ref{Float64}(0.0)
    ptrs, lds = [c.ptr for c in tt.cores], Int64[c.ld for c in tt.cores]
    check(tt.h.ptr, ccall((:tci_tt_sum_f64, LIB), Cint, (Ptr{Cvoid}, Ptr{Ptr{Float64}}, Ptr{Int64}, Ptr{Int64}, Cint, Ref{Float64}),
                          tt.h.ptr, ptrs, tt.shapes, lds, length(tt.cores), out))
    return out[]
end
function evaluate(tt::DeviceTT, points::Vector{MultiIndex})
    n = length(tt.cores)
    idx = Int32[p[b] for b in 1:n, p in points]
    out = Vector{Float64}(undef, length(points))
    ptrs, lds = [c.ptr for c in tt.cores], Int64[c.ld for c in tt.cores]
    check(tt.h.ptr, ccall((:tci_tt_evaluate_f64, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Ptr{Float64}}, Ptr{Int64}, Ptr{Int64}, Cint, Ptr{Int32}, Int64, Ptr{Float64}),
        tt.h.ptr, ptrs, tt.shapes, lds, n, idx, length(points), out))
    return out
end

# ---- optional: fewer host round trips in a sweep ----------------------------------------------------------------------------
# updatemaxsample! (src/tensorci2.jl:636-638) merged on the device: a sweep passes the accumulator to evaluate_device_acc! for
# every Pi and reads it once at the end (the value is only consumed by the tolerance of the next sweep, :1089-1090).
mutable struct MaxAccumulator
    h::Handle
    ptr::Ptr{Cvoid}
    function MaxAccumulator(value::Float64=0.0; h::Handle=handle())
        p = Ref{Ptr{Cvoid}}(C_NULL)
        check(h.ptr, ccall((:tci_malloc, LIB), Cint, (Ptr{Cvoid}, Ref{Ptr{Cvoid}}, Csize_t), h.ptr, p, 64))
        a = new(h, p[])
        finalizer(x -> ccall((:tci_free, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}), x.h.ptr, x.ptr), a)
        check(h.ptr, ccall((:tci_maxacc_set, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Cdouble), h.ptr, a.ptr, value))
        return a
    end
end
function Base.getindex(a::MaxAccumulator)::Float64      # synchronises; max(|initial|, every value since), NaN-propagating
    v = Ref(0.0)
    check(a.h.ptr, ccall((:tci_maxacc_get, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ref{Float64}), a.h.ptr, a.ptr, v))
    return v[]
end
function evaluate_device_acc!(f::B200Function, out::DeviceMatrix, left::Vector{MultiIndex}, right::Vector{MultiIndex},
                              stride_c::Int, stride_j::Int, acc::MaxAccumulator)
    L, R = _table(left), _table(right)
    check(f.h.ptr, ccall((:tci_batcheval_acc_f64, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Int32}, Cint, Ptr{Int32}, Int64, Cint, Ptr{Int32}, Int64, Cint, Ptr{Float64}, Int64, Int64, Ptr{Cvoid}),
        f.h.ptr, f.f, f.localdims, length(f.localdims), L, length(left), size(L, 1), R, length(right), size(R, 1),
        out.ptr, stride_c, stride_j, acc.ptr))
    return out
end

# optfirstpivot(f, localdims, firstpivot; maxsweep) (src/util.jl:260-291) as one call for device functors
function TCI.optfirstpivot(f::B200Function, localdims::Union{Vector{Int},NTuple{N,Int}},
                           firstpivot::MultiIndex=ones(Int, length(localdims)); maxsweep::Int=1000) where {N}
    pivot = Int32.(firstpivot)
    check(f.h.ptr, ccall((:tci_optfirstpivot_f64, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Int32}, Cint, Ptr{Int32}, Cint),
        f.h.ptr, f.f, f.localdims, length(f.localdims), pivot, maxsweep))
    return Int.(pivot)
end

# ---- B3: the solve of setsitetensor! and _contract ------------------------------------------------------------------------
# setsitetensor!(tci, f, b) (src/tensorci2.jl:599-629): Pi1 and P are evaluated on the device and T = Pi1 * inv(P) (:626,
# `transpose(transpose(P) \ transpose(Pi1))`) is tci_solve_right_f64; only the finished site tensor visits the host.
function TCI.setsitetensor!(tci::TCI.TensorCI2{Float64}, f::B200Function, b::Int; leftorthogonal=true)
    leftorthogonal || error("leftorthogonal==false is not supported!")
    Is = leftorthogonal ? TCI.kronecker(tci.Iset[b], tci.localdims[b]) : tci.Iset[b]
    Js = leftorthogonal ? tci.Jset[b] : TCI.kronecker(tci.localdims[b], tci.Jset[b])
    mx = Ref(tci.maxsamplevalue)
    Pi1 = DeviceMatrix(f.h, length(Is), length(Js))
    evaluate_device!(f, Pi1, Is, Js, length(Is), Pi1.ld, mx)                          # :617-620 (+ updatemaxsample!)
    if (leftorthogonal && b == length(tci)) || (!leftorthogonal && b == 1)            # :612-615
        tci.maxsamplevalue = mx[]
        TCI.setsitetensor!(tci, b, Array(Pi1))
        return tci.sitetensors[b]
    end
    P = DeviceMatrix(f.h, length(tci.Iset[b+1]), length(tci.Jset[b]))
    evaluate_device!(f, P, tci.Iset[b+1], tci.Jset[b], length(tci.Iset[b+1]), P.ld, mx)   # :622
    tci.maxsamplevalue = mx[]
    length(tci.Iset[b+1]) == length(tci.Jset[b]) || error("Pivot matrix at bond $(b) is not square!")   # :623
    T = DeviceMatrix(f.h, length(Is), length(Js))
    check(f.h.ptr, ccall((:tci_solve_right_f64, LIB), Cint,
        (Ptr{Cvoid}, Ptr{Float64}, Int64, Int64, Ptr{Float64}, Int64, Int64, Ptr{Float64}, Int64),
        f.h.ptr, Pi1.ptr, Pi1.m, Pi1.ld, P.ptr, P.m, P.ld, T.ptr, T.ld))             # :626
    tci.sitetensors[b] = reshape(Array(T), length(tci.Iset[b]), tci.localdims[b], length(tci.Jset[b]))
    return tci.sitetensors[b]
end

# _contract(a, b, idx_a, idx_b) (src/contraction.jl:193-215) on device tensors: the permutedims / reshape of :211-212 are
# element strides of tci_gemm_f64 (one GEMM, no copies).  Handles the case the contraction code uses everywhere: the
# contracted legs are a contiguous trailing block of `a` and a contiguous leading block of `b`, in the same order
# (idx_a = (k+1, ..., na), idx_b = (1, ..., nb')); anything else goes through tci_strided_copy_f64 first.
struct DeviceTensor{N}
    M::DeviceMatrix                      # dense column-major storage, M.ld == M.m
    dims::NTuple{N,Int}
end
function DeviceTensor(h::Handle, A::Array{Float64,N}) where {N}
    M = DeviceMatrix(h, max(length(A), 1), 1)
    check(h.ptr, ccall((:tci_memcpy_h2d, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Csize_t), h.ptr, M.ptr, A, 8 * length(A)))
    return DeviceTensor{N}(M, size(A))
end
function Base.Array(T::DeviceTensor{N}) where {N}
    out = Array{Float64,N}(undef, T.dims...)
    check(T.M.h.ptr, ccall((:tci_memcpy_d2h, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Csize_t), T.M.h.ptr, out, T.M.ptr, 8 * length(out)))
    return out
end
function TCI._contract(a::DeviceTensor{N1}, b::DeviceTensor{N2}, idx_a::NTuple{n1,Int}, idx_b::NTuple{n2,Int}) where {N1,N2,n1,n2}
    length(idx_a) == length(idx_b) || error("length(idx_a) != length(idx_b)")                       # :200
    length(unique(idx_a)) == length(idx_a) && length(unique(idx_b)) == length(idx_b) || error("idx_a or idx_b contains duplicate elements")
    all(1 <= i <= N1 for i in idx_a) && all(1 <= i <= N2 for i in idx_b) || error("idx_a or idx_b contains elements out of range")
    k = length(idx_a)
    idx_a == ntuple(i -> N1 - k + i, k) && idx_b == ntuple(i -> i, k) || error("TCIB200._contract: permute the operands with tci_strided_copy_f64 first")
    rest_a, rest_b = a.dims[1:N1-k], b.dims[k+1:N2]
    Mr, Kc, Nc = prod(rest_a; init=1), prod(a.dims[N1-k+1:N1]; init=1), prod(rest_b; init=1)
    Kc == prod(b.dims[1:k]; init=1) || error("contracted dimensions do not match")
    h = a.M.h
    C = DeviceMatrix(h, max(Mr * Nc, 1), 1)
    check(h.ptr, ccall((:tci_gemm_f64, LIB), Cint,
        (Ptr{Cvoid}, Int64, Int64, Int64, Float64, Ptr{Float64}, Int64, Int64, Ptr{Float64}, Int64, Int64, Float64, Ptr{Float64}, Int64, Int64,
         Int64, Int64, Int64, Int64),
        h.ptr, Mr, Nc, Kc, 1.0, a.M.ptr, 1, Mr, b.M.ptr, 1, Kc, 0.0, C.ptr, 1, Mr, 1, 0, 0, 0))                    # amat * bmat, :214
    return DeviceTensor{N1 + N2 - 2k}(C, (rest_a..., rest_b...))
end

# ---- the bond loop on the device (csrc/sweep.cu): a whole half-sweep / fillsitetensors! / final sweep1site! per call ----------
# pack Iset / Jset (and the history entries used as extras) into the int32 arena of tci_dsweep_* (include/tci_b200.h)
function _arena(tci::TCI.TensorCI2{Float64}, forward::Bool, s0::Int, s1::Int, maxbonddim::Int,
                extraI::Union{Nothing,Vector{Vector{MultiIndex}}}, extraJ::Union{Nothing,Vector{Vector{MultiIndex}}})
    N = length(tci)
    a = zeros(Int32, 8 + N * 13)
    a[1] = N; a[2] = forward; a[3] = s0; a[4] = s1
    a[5] = maxbonddim >= typemax(Int32) ? 0 : maxbonddim; a[6] = extraI !== nothing
    a[9:8+N] .= tci.localdims
    rows = 0
    for b in 1:N
        tabs = (tci.Iset[b], tci.Jset[b], extraI === nothing ? MultiIndex[] : extraI[b], extraJ === nothing ? MultiIndex[] : extraJ[b])
        for (w, t) in enumerate(tabs)
            d = 8 + N + 12 * (b - 1)
            a[d+2w-1] = length(t); a[d+2w] = length(a)                      # count, word offset (0-based) of the table
            a[d+8+w] = rows; rows += length(t)
            for r in t; append!(a, Int32.(r)); end
        end
    end
    a[7] = rows
    return a
end

"""One half-sweep of sweep2site! (src/tensorci2.jl:1226-1250) as ONE launch wherever the Pi matrices fit shared memory; bonds that
do not fit run through `TCI.updatepivots!` above.  `extraI/extraJ`: the history entries (tensorci2.jl:1212-1218) or nothing."""
function halfsweep!(tci::TCI.TensorCI2{Float64}, f::B200Function, forward::Bool; abstol::Float64, maxbonddim::Int=typemax(Int),
                    extraI=nothing, extraJ=nothing, acc::MaxAccumulator=MaxAccumulator(tci.maxsamplevalue; h=f.h))
    N = length(tci)
    TCI.invalidatesitetensors!(tci)
    s = 0
    while s < N - 1
        a = _arena(tci, forward, s, N - 1, maxbonddim, extraI, extraJ)
        out = Vector{Int32}(undef, 8 + 6 * (N - 1) + (N - 1) * (512 * (N + 1) + 1028))
        used = Ref{Int64}(0)
        check(f.h.ptr, ccall((:tci_dsweep_halfsweep_f64, LIB), Cint,
            (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Int32}, Int64, Float64, Float64, Ptr{Cvoid}, Ptr{Int32}, Int64, Ref{Int64}),
            f.h.ptr, f.f, a, length(a), 1e-14, abstol, acc.ptr, out, length(out), used))
        status, next = out[1], Int(out[2])
        for st in s:next-1
            b = forward ? st + 1 : N - 1 - st                                     # 1-based bond
            rec = 8 + 6 * (st - s)
            r, offA, offB, offE = Int(out[rec+1]), Int(out[rec+2]), Int(out[rec+3]), Int(out[rec+4])
            tci.Iset[b+1] = [Int.(out[offA+1+(k-1)*b:offA+k*b]) for k in 1:r]
            tci.Jset[b] = [Int.(out[offB+1+(k-1)*(N-b):offB+k*(N-b)]) for k in 1:r]
            errs = collect(reinterpret(Float64, out[offE+1:offE+2*(r+1)]))
            TCI.updateerrors!(tci, b, errs)
        end
        status == 2 && error("lu.L contains NaNs")
        status == 3 && error("lu.U contains NaNs")
        s = next
        if status == 1 && s < N - 1                                               # this bond's Pi does not fit: per-bond path
            b = forward ? s + 1 : N - 1 - s
            TCI.updatepivots!(tci, b, f, forward; abstol=abstol, maxbonddim=maxbonddim,
                              extraIset=extraI === nothing ? MultiIndex[] : extraI[b+1], extraJset=extraJ === nothing ? MultiIndex[] : extraJ[b])
            s += 1
        end
    end
    tci.maxsamplevalue = acc[]
    nothing
end

"""fillsitetensors!(tci, f) (src/globalsearch.jl:202-208) for all sites in one launch (bond dimensions <= 64)."""
function TCI.fillsitetensors!(tci::TCI.TensorCI2{Float64}, f::B200Function)
    N = length(tci)
    if maximum(length.(tci.Jset[1:N-1])) > 64
        for b in 1:N; TCI.setsitetensor!(tci, f, b); end
        return nothing
    end
    acc = MaxAccumulator(tci.maxsamplevalue; h=f.h)
    cores = [DeviceMatrix(f.h, length(tci.Iset[b]) * tci.localdims[b], length(tci.Jset[b])) for b in 1:N]
    a = _arena(tci, true, 0, 0, typemax(Int), nothing, nothing)
    ptrs = [c.ptr for c in cores]
    # (cores are allocated with ld == rows here: DeviceMatrix pads ld, so request dense buffers)
    check(f.h.ptr, ccall((:tci_dsweep_fill_f64, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Int32}, Int64, Ptr{Cvoid}, Ptr{Ptr{Float64}}),
                         f.h.ptr, f.f, a, length(a), acc.ptr, ptrs))
    for b in 1:N
        q, r = length(tci.Iset[b]) * tci.localdims[b], length(tci.Jset[b])
        host = Vector{Float64}(undef, q * r)
        check(f.h.ptr, ccall((:tci_memcpy_d2h, LIB), Cint, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Csize_t), f.h.ptr, host, cores[b].ptr, 8 * q * r))
        tci.sitetensors[b] = reshape(host, length(tci.Iset[b]), tci.localdims[b], r)
    end
    tci.maxsamplevalue = acc[]
    nothing
end

end # module
