# B200Ext.jl -- the package extension a DINCAE.jl maintainer would add to bind libdincae_b200.so.
#
# NOT EXECUTED IN THIS REPOSITORY: the build image has no Julia runtime (and no network to install one), so this
# file is written by inspection against include/dincae_b200.h.  The Python ctypes binding in dincae.jl_b200/lib.py
# drives the very same symbols in the same order and is what the test-suite runs.
#
# Usage from DINCAE:   DINCAE.reconstruct(B200Ext.B200Array{Float32}, data_all, fnames_rec; kw...)
# The reference's epoch loop (src/model.jl:588-651) stays as it is; only the methods below are new.
module B200Ext

import DINCAE: interpnd!, interp_adjn!, _to_device, train_init, train_epoch!, loss_function, genmodel

const lib = "libdincae_b200"
const HOST = Int32(0)

"Tag type passed as `Atype` (plays the role of `CuArray{Float32}`); host arrays stay plain `Array`s."
struct B200Array{T,N} <: AbstractArray{T,N} end

# ---- error convention: non-zero status -> Julia exception -------------------------------------------------------
check(st::Integer) = st == 0 || error(unsafe_string(ccall((:dincae_last_error, lib), Cstring, ())))

# ---- handles ------------------------------------------------------------------------------------------------------
mutable struct Ctx
    h::Ptr{Cvoid}
    function Ctx(device::Integer = 0)
        r = Ref{Ptr{Cvoid}}(C_NULL)
        check(ccall((:dincae_ctx_create, lib), Int32, (Int32, Ref{Ptr{Cvoid}}), device, r))
        c = new(r[])
        finalizer(c -> ccall((:dincae_ctx_destroy, lib), Int32, (Ptr{Cvoid},), c.h), c)
    end
end
const default_ctx = Ref{Union{Nothing,Ctx}}(nothing)
ctx() = something(default_ctx[], (default_ctx[] = Ctx(); default_ctx[]))

# layout-compatible with `dincae_config` (include/dincae_b200.h): 35 4-byte fields, no padding
struct Config
    W::Int32; H::Int32; ninput::Int32; noutput::Int32; nlevels::Int32
    nfilter::NTuple{8,Int32}
    skip_mask::UInt32
    upsampling::Int32
    nchains::Int32
    loss_weights::NTuple{4,Float32}
    gamma::Float32
    truth_uncertain::Int32
    l1_beta::Float32; l2_beta::Float32
    laplacian_penalty::Float32; laplacian_error_penalty::Float32
    clip_grad::Float32
    adam_beta1::Float32; adam_beta2::Float32; adam_eps::Float32
    leaky_slope::Float32
    precision::Int32
    max_batch::Int32
end

pad8(v) = ntuple(i -> i <= length(v) ? Int32(v[i]) : Int32(0), 8)
pad4(v) = ntuple(i -> i <= length(v) ? Float32(v[i]) : 0f0, 4)

mutable struct Model
    h::Ptr{Cvoid}
    ctx::Ctx
    cfg::Config
end

"genmodel(...) |> device + train_init (src/model.jl:546-559, :585): same keywords as DINCAE.genmodel"
function genmodel(::Type{<:B200Array}, sz, ninput, noutput; batch_size,
                  truth_uncertain = false, enc_nfilter_internal = [16,24,36,54],
                  skipconnections = 2:(length(enc_nfilter_internal)+1),
                  regularization_L1_beta = 0, regularization_L2_beta = 0, upsampling_method = :nearest,
                  loss_weights_refine = (1.,), min_std_err = exp(-5), laplacian_penalty = 0,
                  laplacian_error_penalty = laplacian_penalty, clip_grad = 5.0, precision = :tf32)
    L = length(enc_nfilter_internal)
    mask = UInt32(0)
    for s in skipconnections
        1 <= s <= L && (mask |= UInt32(1) << (s-1))      # level L+1 is the identity (src/model.jl:263-264)
    end
    cfg = Config(sz[1], sz[2], ninput, noutput, L, pad8(enc_nfilter_internal), mask,
                 upsampling_method == :nearest ? 0 : 1, length(loss_weights_refine), pad4(loss_weights_refine),
                 Float32(log(min_std_err^(-2))), truth_uncertain, regularization_L1_beta, regularization_L2_beta,
                 laplacian_penalty, laplacian_error_penalty, isnothing(clip_grad) ? 0f0 : clip_grad,
                 0.9f0, 0.999f0, 1f-8, 0f0, precision == :fp32 ? 0 : 1, batch_size)
    r = Ref{Ptr{Cvoid}}(C_NULL)
    c = ctx()
    check(ccall((:dincae_model_create, lib), Int32, (Ptr{Cvoid}, Ref{Config}, Ref{Ptr{Cvoid}}), c.h, Ref(cfg), r))
    m = Model(r[], c, cfg)
    finalizer(m -> ccall((:dincae_model_destroy, lib), Int32, (Ptr{Cvoid},), m.h), m)
    check(ccall((:dincae_model_init_glorot, lib), Int32, (Ptr{Cvoid}, UInt64), m.h, rand(UInt64)))
    return m
end

"model(xin): (W,H,ninput,B) -> (W,H,2*noutput,B) = (mean, sigma^2)  (StepModel call, src/model.jl:182-189)"
function (m::Model)(xin::Array{Float32,4})
    out = Array{Float32}(undef, size(xin,1), size(xin,2), 2*m.cfg.noutput, size(xin,4))
    check(ccall((:dincae_forward, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Int32, Ptr{Float32}, Int32),
                m.h, xin, size(xin,4), out, HOST))
    return out
end

"loss_function(model, xin, xtrue) (src/model.jl:192-222)"
function loss_function(m::Model, xin::Array{Float32,4}, xtrue::Array{Float32,4})
    l = Ref{Float64}()
    check(ccall((:dincae_loss, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Int32, Int32, Ref{Float64}),
                m.h, xin, xtrue, size(xin,4), HOST, l))
    return l[]
end

train_init(m::Model, optim; clip_grad = nothing, learning_rate = nothing) = (m, nothing, nothing)   # Adam state lives in the library

"train_epoch! (src/flux.jl:24-40): mean of the per-batch losses; the next batch is uploaded while this one trains"
function train_epoch!((m, _, _)::Tuple{Model,Nothing,Nothing}, dl, learning_rate; clip_grad = nothing)
    loss_sum = 0.0; N = 0; loss = Ref{Float32}()
    it = iterate(dl)
    it === nothing && error("empty data iterator")
    (xin, xtrue), state = it
    slot = Int32(0)
    check(ccall((:dincae_batch_upload, lib), Int32, (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Int32), m.h, slot, xin, xtrue, size(xin,4)))
    while true
        nxt = iterate(dl, state)
        if nxt !== nothing
            (xin2, xtrue2), state = nxt
            check(ccall((:dincae_batch_upload, lib), Int32, (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Int32), m.h, slot ⊻ Int32(1), xin2, xtrue2, size(xin2,4)))
        end
        GC.@preserve xin xtrue begin
            check(ccall((:dincae_train_step_slot, lib), Int32, (Ptr{Cvoid}, Int32, Int32, Float32, Ref{Float32}), m.h, slot, size(xin,4), learning_rate, loss))
        end
        loss_sum += loss[]; N += 1
        nxt === nothing && break
        xin, xtrue, slot = xin2, xtrue2, slot ⊻ Int32(1)
    end
    return loss_sum / N
end

# ---- the two kernels of ext/CUDAExt.jl ------------------------------------------------------------------------------
"interpnd!(pos, A, vec) (ext/CUDAExt.jl:31): NTuple{2,Float32} positions are interleaved (x,y) pairs in memory"
function interpnd!(pos::Vector{NTuple{2,Float32}}, A::Matrix{Float32}, vec::Vector{Float32}, ::Type{<:B200Array})
    check(ccall((:dincae_interp, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Int64, Ptr{Float32}, Int32, Int32, Ptr{Float32}, Int32),
                ctx().h, pos, length(pos), A, size(A,1), size(A,2), vec, HOST))
    return nothing
end

"interp_adjn!(pos, values, B) (ext/CUDAExt.jl:71): deterministic, no float atomics"
function interp_adjn!(pos::Vector{NTuple{2,Float32}}, values::Vector{Float32}, B::Matrix{Float32}, ::Type{<:B200Array})
    check(ccall((:dincae_interp_adjoint, lib), Int32, (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Int64, Int32, Int32, Ptr{Float32}, Int32),
                ctx().h, pos, values, length(pos), size(B,1), size(B,2), B, HOST))
    return nothing
end

_to_device(::Type{<:B200Array}) = identity     # data stays in host Arrays at this boundary; the library stages it

end # module
