// Host-side CUtensorMap construction.  cuTensorMapEncodeTiled is fetched through the runtime's driver
// entry-point query, so the library has no link-time dependency on libcuda.
#include "umma.cuh"
#include "common.cuh"

#include <cudaTypedefs.h>

namespace bsa {

namespace {
PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (PFN_cuTensorMapEncodeTiled_v12000)p;
    }
    return fn;
}
}  // namespace

int make_tmap_2d(CUtensorMap* map, const float* base, long long rows, long long cols, long long pitch_elems, int box_rows,
                 bool atom32) {
    // The driver entry point needs a current context on the calling thread.  A thread that has made no runtime call yet
    // (PyTorch's autograd worker entering the backward pass with every buffer served from the caching allocator) has
    // none bound and the encode fails with CUDA_ERROR_INVALID_CONTEXT; cudaFree(0) binds the primary context.
    static thread_local bool ctx_bound = false;
    if (!ctx_bound) {
        BSA_CHECK_CUDA(cudaFree(nullptr));
        ctx_bound = true;
    }
    auto enc = get_encode();
    BSA_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
    BSA_REQUIRE(((uintptr_t)base & 15) == 0 && (pitch_elems * 4) % 16 == 0,
                "TMA operand must be 16-byte aligned with a 16-byte-multiple row pitch (pitch %lld elems)", pitch_elems);
    BSA_REQUIRE(box_rows >= 1 && box_rows <= 256, "TMA box rows %d out of range", box_rows);
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch_elems * sizeof(float)};
    cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BSA_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rows %lld cols %lld pitch %lld box %d)",
                (int)r, rows, cols, pitch_elems, box_rows);
    return 0;
}

}  // namespace bsa
