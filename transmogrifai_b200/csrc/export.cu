// OPVector -> Spark ml.linalg VectorUDT, columnar: what a mapPartitions transform hands back to Spark for the rows of a
// partition.  The reference ends every vectorizer / combiner / SanityCheckerModel with `.compressed`
// (Transmogrifier.scala:466-475, RichVector.scala:120-134, DerivedFeatureFilterUtils.scala:291-305): explicit zeros are
// dropped and a row becomes a SparseVector when 1.5 (nnz + 1) < size, a DenseVector otherwise (Spark 2.4.5
// ml.linalg.Vector.compressed).  VectorUDT's struct is (type: byte 0 sparse / 1 dense, size: int -- null for dense,
// indices: int[] -- null for dense, values: double[]); as Arrow arrays of a partition: type u8 [n], size i32 [n],
// indices = offsets i64 [n + 1] + i32 [], values = offsets i64 [n + 1] + f64 [].
//
// Two HBM-bound kernels, a warp per row: count (non-zeros -> type and the two list lengths), then -- after a prefix sum
// of the lengths -- write.  The parts of the vector are read where the vectorizers left them (dense blocks, one-hot
// index blocks, sparse rows in 4-array form); nothing is densified.
#include <algorithm>

#include "common.cuh"

namespace tmog {

struct ExportPart {        // device struct, 64 bytes
    int64_t kind;          // 0 dense f64 [n][ld], 1 one-hot i32 [n][ld] (k entries, -1 = none), 2 sparse rows
    int64_t width;         // columns of the part
    int64_t col0;          // first vector index of the part
    int64_t ld;            // row stride (kind 0, 1); entries per row k (kind 1) in `k`
    int64_t k;
    const void* a;         // kind 0: values; kind 1: indices; kind 2: row_start (i64)
    const void* b;         // kind 2: row_nnz (i32)
    const void* c;         // kind 2: indices (i32)
    const void* d;         // kind 2: values (i32)
};

template <bool WRITE>
__global__ void __launch_bounds__(256)
vector_export_kernel(const ExportPart* __restrict__ parts, int nparts, int64_t n, int size, uint8_t* __restrict__ type,
                     int32_t* __restrict__ size_out, int64_t* __restrict__ idx_len, int64_t* __restrict__ val_len,
                     const int64_t* __restrict__ idx_off, const int64_t* __restrict__ val_off,
                     int32_t* __restrict__ indices, double* __restrict__ values) {
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n; r += warps) {
        bool sparse = false;
        int64_t io = 0, vo = 0;
        if (WRITE) {
            sparse = type[r] == 0;
            io = idx_off[r];
            vo = val_off[r];
            if (!sparse)
                for (int q = lane; q < size; q += 32) values[vo + q] = 0.0;
            __syncwarp();
        }
        int nnz = 0;
        for (int p = 0; p < nparts; ++p) {
            const ExportPart e = parts[p];
            if (e.kind == 0) {
                const double* x = static_cast<const double*>(e.a) + r * e.ld;
                for (int64_t c0 = 0; c0 < e.width; c0 += 32) {
                    const int64_t c = c0 + lane;
                    const double v = c < e.width ? x[c] : 0.0;
                    const unsigned nz = __ballot_sync(FULL, v != 0.0);
                    if (WRITE && v != 0.0) {
                        if (sparse) {
                            const int64_t at = nnz + __popc(nz & lt);
                            indices[io + at] = (int32_t)(e.col0 + c);
                            values[vo + at] = v;
                        } else {
                            values[vo + e.col0 + c] = v;
                        }
                    }
                    nnz += __popc(nz);
                }
            } else if (e.kind == 1) {
                const int32_t* x = static_cast<const int32_t*>(e.a) + r * e.ld;
                for (int64_t c0 = 0; c0 < e.k; c0 += 32) {
                    const int64_t c = c0 + lane;
                    const int32_t v = c < e.k ? x[c] : -1;
                    const unsigned nz = __ballot_sync(FULL, v >= 0);
                    if (WRITE && v >= 0) {
                        if (sparse) {
                            const int64_t at = nnz + __popc(nz & lt);
                            indices[io + at] = (int32_t)(e.col0 + v);
                            values[vo + at] = 1.0;
                        } else {
                            values[vo + e.col0 + v] = 1.0;
                        }
                    }
                    nnz += __popc(nz);
                }
            } else {
                const int64_t s = static_cast<const int64_t*>(e.a)[r];
                const int m = static_cast<const int32_t*>(e.b)[r];
                const int32_t* ix = static_cast<const int32_t*>(e.c) + s;
                const int32_t* vx = static_cast<const int32_t*>(e.d) + s;
                for (int c0 = 0; c0 < m; c0 += 32) {
                    const int c = c0 + lane;
                    const int32_t v = c < m ? vx[c] : 0;
                    const unsigned nz = __ballot_sync(FULL, v != 0);
                    if (WRITE && v != 0) {
                        if (sparse) {
                            const int64_t at = nnz + __popc(nz & lt);
                            indices[io + at] = (int32_t)(e.col0 + ix[c]);
                            values[vo + at] = (double)v;
                        } else {
                            values[vo + e.col0 + ix[c]] = (double)v;
                        }
                    }
                    nnz += __popc(nz);
                }
            }
        }
        if (!WRITE && lane == 0) {
            const bool sp = 1.5 * (nnz + 1.0) < (double)size;         // Vector.compressed
            type[r] = sp ? 0 : 1;
            size_out[r] = sp ? size : -1;                              // -1 = null (dense rows carry no size)
            idx_len[r] = sp ? nnz : 0;
            val_len[r] = sp ? nnz : size;
        }
    }
}

// ---- compact return of a chunk's vector over PCIe (streaming.HostSink(compact=True)): what the device-to-host copy of a
//      transformed partition carries.  Half of a RealVectorizer / BinaryVectorizer block is null indicators (0.0 / 1.0 as
//      f64, RealVectorizer.scala / OpScalarStandardScaler aside), and a hashed text entry's term frequency is a small
//      integer; both travel as one byte.
// dense f64 [n][ld] -> values f64 [n][nv] (columns with unit_col == 0, in order) + units u8 [n][nu] (columns with
// unit_col == 1: 0 / 1 valued by construction); flags[0] |= 1 if a unit column holds anything else
__global__ void __launch_bounds__(256)
dense_split_kernel(const double* __restrict__ data, int64_t ld, int64_t n, int ncols, const int32_t* __restrict__ dest,
                   double* __restrict__ values, int64_t ldv, uint8_t* __restrict__ units, int64_t ldu,
                   int32_t* __restrict__ flags) {
    // dest[c] >= 0: value column number dest[c]; dest[c] < 0: unit column number ~dest[c]
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    bool bad = false;
    for (int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n; r += warps) {
        const double* x = data + r * ld;
        for (int c = lane; c < ncols; c += 32) {
            const double v = x[c];
            const int d = dest[c];
            if (d >= 0) {
                values[r * ldv + d] = v;
            } else {
                bad |= !(v == 0.0 || v == 1.0);
                units[r * ldu + ~d] = v != 0.0 ? 1 : 0;
            }
        }
    }
    if (bad) flags[0] = 1;
}

// int32 column indices of a block of at most 65 536 columns -> u16 (eight per thread and step)
__global__ void __launch_bounds__(256)
indices_narrow_kernel(const int32_t* __restrict__ idx, int64_t n, const unsigned long long* __restrict__ used,
                      unsigned short* __restrict__ out, int32_t* __restrict__ flags) {
    if (used) n = min(n, (int64_t)*used);          // entries behind the last one written are not looked at
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 8;
    bool bad = false;
    for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i0 < n; i0 += stride) {
        if (i0 + 8 <= n) {
            const int4 a = *reinterpret_cast<const int4*>(idx + i0), b = *reinterpret_cast<const int4*>(idx + i0 + 4);
            bad |= ((unsigned)(a.x | a.y | a.z | a.w | b.x | b.y | b.z | b.w) >> 16) != 0u;
            uint4 o;
            o.x = (unsigned)(a.x & 0xffff) | ((unsigned)a.y << 16);
            o.y = (unsigned)(a.z & 0xffff) | ((unsigned)a.w << 16);
            o.z = (unsigned)(b.x & 0xffff) | ((unsigned)b.y << 16);
            o.w = (unsigned)(b.z & 0xffff) | ((unsigned)b.w << 16);
            *reinterpret_cast<uint4*>(out + i0) = o;
        } else {
            for (int64_t i = i0; i < n; ++i) {
                bad |= ((unsigned)idx[i] >> 16) != 0u;
                out[i] = (unsigned short)idx[i];
            }
        }
    }
    if (bad) flags[1] = 1;
}

// int32 values -> u8; a value above 254 becomes 255 and goes to the overflow list (position, value) as well
__global__ void __launch_bounds__(256)
values_narrow_kernel(const int32_t* __restrict__ values, int64_t n, const unsigned long long* __restrict__ used,
                     uint8_t* __restrict__ out, int64_t* __restrict__ over_pos, int32_t* __restrict__ over_val, int over_cap,
                     unsigned* __restrict__ over_count) {
    if (used) n = min(n, (int64_t)*used);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i0 < n; i0 += stride) {
        if (i0 + 4 <= n) {
            const int4 v = *reinterpret_cast<const int4*>(values + i0);
            const int w[4] = {v.x, v.y, v.z, v.w};
            unsigned packed = 0u;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const bool big = w[q] > 254 || w[q] < 0;
                packed |= (big ? 255u : (unsigned)w[q]) << (8 * q);
                if (big) {
                    const unsigned k = atomicAdd(over_count, 1u);
                    if (k < (unsigned)over_cap) {
                        over_pos[k] = i0 + q;
                        over_val[k] = w[q];
                    }
                }
            }
            *reinterpret_cast<unsigned*>(out + i0) = packed;
        } else {
            for (int64_t i = i0; i < n; ++i) {
                const int w = values[i];
                const bool big = w > 254 || w < 0;
                out[i] = big ? 255 : (uint8_t)w;
                if (big) {
                    const unsigned k = atomicAdd(over_count, 1u);
                    if (k < (unsigned)over_cap) {
                        over_pos[k] = i;
                        over_val[k] = w;
                    }
                }
            }
        }
    }
}

}  // namespace tmog

using namespace tmog;

extern "C" int tmog_vector_export_count_dev(const void* parts, int nparts, int64_t n, int size, uint8_t* type,
                                            int32_t* size_out, int64_t* idx_len, int64_t* val_len, void* stream) {
    TMOG_REQUIRE(nparts >= 0 && n >= 0 && size >= 0, "bad shape");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(parts && type && size_out && idx_len && val_len, "NULL argument");
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 7) / 8, (int64_t)sm_count() * 16));
    vector_export_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(static_cast<const ExportPart*>(parts), nparts, n,
                                                                        size, type, size_out, idx_len, val_len, nullptr,
                                                                        nullptr, nullptr, nullptr);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_vector_export_write_dev(const void* parts, int nparts, int64_t n, int size, const uint8_t* type,
                                            const int64_t* idx_off, const int64_t* val_off, int32_t* indices,
                                            double* values, void* stream) {
    TMOG_REQUIRE(nparts >= 0 && n >= 0 && size >= 0, "bad shape");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(parts && type && idx_off && val_off && indices && values, "NULL argument");
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 7) / 8, (int64_t)sm_count() * 16));
    vector_export_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(
        static_cast<const ExportPart*>(parts), nparts, n, size, const_cast<uint8_t*>(type), nullptr, nullptr, nullptr,
        idx_off, val_off, indices, values);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_dense_split_dev(const double* data, int64_t ld, int64_t n, int ncols, const int32_t* dest,
                                    double* values, int64_t ldv, uint8_t* units, int64_t ldu, int32_t* flags, void* stream) {
    TMOG_REQUIRE(n >= 0 && ncols >= 0 && ld >= ncols, "bad shape");
    if (n == 0 || ncols == 0) return TMOG_OK;
    TMOG_REQUIRE(data && dest && flags && (values || ldv == 0) && (units || ldu == 0), "NULL argument");
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 7) / 8, (int64_t)sm_count() * 16));
    dense_split_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(data, ld, n, ncols, dest, values, ldv, units, ldu, flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_indices_narrow_dev(const int32_t* indices, int64_t n, const uint64_t* used, uint16_t* out, int32_t* flags,
                                       void* stream) {
    TMOG_REQUIRE(n >= 0, "bad shape");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(indices && out && flags, "NULL argument");
    TMOG_REQUIRE((reinterpret_cast<uintptr_t>(indices) & 15u) == 0 && (reinterpret_cast<uintptr_t>(out) & 15u) == 0,
                 "indices and out must be 16-byte aligned");
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n / 8 + 255) / 256, (int64_t)sm_count() * 16));
    indices_narrow_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(indices, n, reinterpret_cast<const unsigned long long*>(used), out,
                                                                 flags);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}

extern "C" int tmog_values_narrow_dev(const int32_t* values, int64_t n, const uint64_t* used, uint8_t* out, int64_t* over_pos,
                                      int32_t* over_val, int over_cap, uint32_t* over_count, void* stream) {
    TMOG_REQUIRE(n >= 0 && over_cap >= 0, "bad shape");
    if (n == 0) return TMOG_OK;
    TMOG_REQUIRE(values && out && over_count && (over_cap == 0 || (over_pos && over_val)), "NULL argument");
    TMOG_REQUIRE((reinterpret_cast<uintptr_t>(values) & 15u) == 0 && (reinterpret_cast<uintptr_t>(out) & 3u) == 0,
                 "values must be 16-byte aligned, out 4-byte aligned");
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n / 4 + 255) / 256, (int64_t)sm_count() * 16));
    values_narrow_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(values, n, reinterpret_cast<const unsigned long long*>(used), out,
                                                                over_pos, over_val, over_cap, over_count);
    TMOG_LAUNCH_CHECK();
    return TMOG_OK;
}
