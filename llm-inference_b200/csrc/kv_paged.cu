// Paged KV cache ("KV-cache layout evolution ... paged blocks").
// The reference allocates K and V as [layers, B, hk, max_seq, d] (src/models/llama/llama.cpp:77-78): every sequence
// slot owns max_seq rows whether it uses them or not. Here a layer's K (V) is a POOL of pages
//     pool [n_pages, hk, B200_KV_PAGE_ROWS = 128 rows, d]
// and sequence b owns the pages listed in page_table[b * max_pages + 0 .. ceil(len_b / 128)) - one list for all layers
// (layer l uses the same page ids in its own pool). A page holds the 128 rows of ALL kv heads of one sequence, so one
// list entry serves the whole (sequence, layer) slice: 1 MB of K per page at 7B (hk = 32, d = 128, fp16), every
// (head, page) run the attention streams is 32 KB contiguous - far above the size where HBM3e reads stop being
// sector-efficient - and 128 rows = the key tile of the tcgen05 context attention and a multiple of the 8 keys one warp
// iteration of the decode attention loads, so a tile never straddles two pages.
// Who reads through the list: decode_attn_split_kernel<..., PAGED> (decode_attention.cu; it also appends the new row).
// The prompt pass attends over a dense fp16 scratch slice and scatters its rows into the pages afterwards (below), the
// same shape the fp8 cache uses (kv_fp8.cu).
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

constexpr int kPageRows = B200_KV_PAGE_ROWS;

// rows t < n_rows[b] of a dense slice src [B, hk, src_seq, d] <-> rows row0[b] + t of sequence b in the pool.
// One thread per 16-byte packet; TO_PAGES: dense -> pool (the prompt pass), else pool -> dense (tests, debugging).
template <bool TO_PAGES>
__global__ void __launch_bounds__(256)
kv_pages_copy_kernel(char* dense, char* pool, const int* __restrict__ page_table, int max_pages,
                     const int* __restrict__ n_rows, const int* __restrict__ row0, int dense_row0, int kv_head_num,
                     int dense_seq, int row_bytes, int max_rows) {
    const int ppr = row_bytes >> 4;   // packets per row
    const int bh = blockIdx.y, b = bh / kv_head_num, kvh = bh - b * kv_head_num;
    const int rows = min(n_rows[b], max_rows);
    const int r0 = row0 ? row0[b] : 0;
    const int* pages = page_table + (size_t)b * max_pages;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows * ppr; i += gridDim.x * blockDim.x) {
        const int t = i / ppr, pk = i - t * ppr;
        const int row = r0 + t;
        if (row < 0 || row >= max_pages * kPageRows) continue;
        const int pg = pages[row / kPageRows];
        char* pp = pool + ((((size_t)pg * kv_head_num + kvh) * kPageRows + (size_t)(row % kPageRows)) * row_bytes) + ((size_t)pk << 4);
        char* dp = dense + (((size_t)bh * dense_seq + (size_t)(dense_row0 + (TO_PAGES ? t : row))) * row_bytes) + ((size_t)pk << 4);
        if (TO_PAGES) *reinterpret_cast<uint4*>(pp) = *reinterpret_cast<const uint4*>(dp);
        else *reinterpret_cast<uint4*>(dp) = *reinterpret_cast<const uint4*>(pp);
    }
}

}  // namespace b200

using namespace b200;

static int pages_copy(bool to_pages, void* dense, void* pool, const int* page_table, int max_pages, const int* n_rows,
                      const int* row0, int batch, int kv_head_num, int dense_seq, int head_size, int esz, int max_rows,
                      void* stream) {
    B200_RETURN_IF(!dense || !pool || !page_table || !n_rows, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || kv_head_num <= 0 || dense_seq <= 0 || head_size <= 0 || max_pages <= 0 || max_rows < 0, B200_EINVAL);
    B200_RETURN_IF(esz != 1 && esz != 2 && esz != 4, B200_EINVAL);
    const long long row_bytes = (long long)head_size * esz;
    B200_RETURN_IF(row_bytes % 16 != 0 || row_bytes > (1 << 20), B200_EUNSUPPORTED);
    B200_RETURN_IF((reinterpret_cast<uintptr_t>(dense) & 15u) || (reinterpret_cast<uintptr_t>(pool) & 15u), B200_EINVAL);
    B200_RETURN_IF((long long)batch * kv_head_num > 65535, B200_EUNSUPPORTED);
    if (max_rows == 0) return 0;
    const long long packets = (long long)max_rows * (row_bytes / 16);
    long long bx = (packets + 255) / 256;
    if (bx > 1024) bx = 1024;
    const dim3 grid((unsigned)bx, (unsigned)(batch * kv_head_num));
    if (to_pages)
        kv_pages_copy_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>((char*)dense, (char*)pool, page_table, max_pages, n_rows,
                                                                            row0, 0, kv_head_num, dense_seq, (int)row_bytes, max_rows);
    else
        kv_pages_copy_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>((char*)dense, (char*)pool, page_table, max_pages, n_rows,
                                                                             row0, 0, kv_head_num, dense_seq, (int)row_bytes, max_rows);
    count_launch();
    return (int)cudaGetLastError();
}

/* rows t < n_rows[b] of the dense slice src [batch, kv_heads, src_seq, d] -> rows dst_row0[b] + t of sequence b's pages
 * (n_rows, dst_row0, page_table: DEVICE int; dst_row0 NULL = 0; max_rows >= every n_rows[b] sizes the launch) */
extern "C" int b200_kv_scatter_pages(const void* src, void* pool, const int* page_table, int max_pages, const int* n_rows,
                                     const int* dst_row0, int batch, int kv_head_num, int src_seq, int head_size,
                                     int elem_bytes, int max_rows, void* stream) {
    return pages_copy(true, const_cast<void*>(src), pool, page_table, max_pages, n_rows, dst_row0, batch, kv_head_num, src_seq,
                      head_size, elem_bytes, max_rows, stream);
}

/* rows [0, n_rows[b]) of sequence b's pages -> rows [0, n_rows[b]) of the dense slice dst [batch, kv_heads, dst_seq, d] */
extern "C" int b200_kv_gather_pages(const void* pool, void* dst, const int* page_table, int max_pages, const int* n_rows,
                                    int batch, int kv_head_num, int dst_seq, int head_size, int elem_bytes, int max_rows,
                                    void* stream) {
    return pages_copy(false, dst, const_cast<void*>(pool), page_table, max_pages, n_rows, nullptr, batch, kv_head_num, dst_seq,
                      head_size, elem_bytes, max_rows, stream);
}
