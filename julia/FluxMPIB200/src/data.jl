# DistributedDataContainer (reference: src/data.jl:8-26): contiguous ceil-sized blocks in rank
# order over anything with `length` / `getindex` (the MLUtils contract). Trailing ranks get an
# empty shard where the reference would throw a BoundsError (N=9, W=4).

"""
    DistributedDataContainer(data)
"""
struct DistributedDataContainer{D}
    data::D
    idxs::UnitRange{Int}
end

function DistributedDataContainer(data)
    total = length(data)
    per = cld(total, total_workers())
    lo = min(local_rank() * per, total)
    hi = min(lo + per, total)
    return DistributedDataContainer(data, (lo + 1):hi)
end

Base.length(ddc::DistributedDataContainer) = length(ddc.idxs)
Base.getindex(ddc::DistributedDataContainer, i) = getindex(ddc.data, ddc.idxs[i])
