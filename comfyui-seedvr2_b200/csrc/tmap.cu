#include "tmap.cuh"

#include "common.cuh"

namespace svr2 {
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

int encode(CUtensorMap* m, const void* ptr, int rank, const cuuint64_t* gdim, const cuuint64_t* gstr, const cuuint32_t* box,
           const cuuint32_t* estr_in = nullptr) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return SVR2_E_CUDA;
    }
    const cuuint32_t ones[4] = {1, 1, 1, 1};
    const cuuint32_t* estr = estr_in ? estr_in : ones;
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), gdim, gstr, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (rank %d, dims %llu x %llu, CUresult %d)", rank,
                  (unsigned long long)gdim[0], (unsigned long long)gdim[rank - 1], (int)r);
        return SVR2_E_CUDA;
    }
    return SVR2_OK;
}

}  // namespace

int make_tmap_box(CUtensorMap* m, const void* ptr, int64_t P, int heads, int box_rows) {
    const cuuint64_t gdim[3] = {(cuuint64_t)SVR2_HEAD_DIM, (cuuint64_t)heads, (cuuint64_t)P};
    const cuuint64_t gstr[2] = {(cuuint64_t)SVR2_HEAD_DIM * 2, (cuuint64_t)heads * SVR2_HEAD_DIM * 2};
    const cuuint32_t box[3] = {64, 1, (cuuint32_t)box_rows};
    return encode(m, ptr, 3, gdim, gstr, box);
}

int make_tmap_2d(CUtensorMap* m, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
    const cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    return encode(m, ptr, 2, gdim, gstr, box);
}

int make_tmap_thwc(CUtensorMap* m, const void* ptr, int64_t T, int64_t H, int64_t W, int64_t C, int box_w, int box_h, int stride) {
    const cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)T};
    const cuuint64_t gstr[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    // with an element stride s the unit loads ceil(boxDim / s) elements: boxDim = N * s gathers N positions, every s-th
    const cuuint32_t box[4] = {64, (cuuint32_t)(box_w * stride), (cuuint32_t)(box_h * stride), 1};
    const cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    return encode(m, ptr, 4, gdim, gstr, box, estr);
}

}  // namespace svr2
