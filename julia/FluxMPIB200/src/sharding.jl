# DistributedDataContainer: rank r keeps the r-th contiguous block of ceil(N / W) samples of any
# container with `length` / `getindex` (contract: /root/reference/src/data.jl:8-26 and
# test/test_data.jl:15-26). Ranks past the end of the data hold an empty range.

struct DistributedDataContainer{D}
    data::D
    idxs::UnitRange{Int}
end

function _block(total::Int, rank::Int, workers::Int)
    width = cld(total, workers)
    first = min(rank * width, total) + 1
    return first:min(first + width - 1, total)
end

DistributedDataContainer(data) =
    DistributedDataContainer(data, _block(length(data), local_rank(), total_workers()))

Base.length(c::DistributedDataContainer) = length(c.idxs)
Base.getindex(c::DistributedDataContainer, i) = c.data[c.idxs[i]]
