/* mtb200_host.h — C ABI of libmtb200_host.so (CPU-side runtime, no CUDA needed).
 *
 * Stand-ins for the native pieces the reference only reaches through its
 * dependencies: the MPI runtime and its collectives (N1-N4)
 * and the IDX/gzip dataset parser (N10).
 */
#ifndef MTB200_HOST_H_
#define MTB200_HOST_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  MTB_OK = 0,
  MTB_ERR_ARG = 1,      /* bad argument */
  MTB_ERR_SYS = 2,      /* OS call failed (errno set) */
  MTB_ERR_TIMEOUT = 3,  /* a peer did not arrive within the timeout */
  MTB_ERR_ABORTED = 4,  /* another rank raised the abort word */
  MTB_ERR_FORMAT = 5    /* malformed input file */
};

/* ---- shared-memory process group (replaces MPI_Init / COMM_WORLD, mpipy.py:5,208-210) */
int mtb_shm_open(const char* name, int rank, int world, uint64_t slot_bytes, int timeout_ms,
                 void** out_handle);
int mtb_shm_close(void* h);
int mtb_shm_rank(void* h);                                  /* comm.Get_rank(), mpipy.py:209 */
int mtb_shm_size(void* h);                                  /* comm.Get_size(), mpipy.py:210 */
void mtb_shm_set_timeout(void* h, int timeout_ms);
void mtb_shm_abort(void* h);                                /* fail-fast: wake every waiter */
uint64_t mtb_shm_heartbeat(void* h, int rank);
int mtb_shm_barrier(void* h);                               /* comm.Barrier(), mpipy.py:93 */
int mtb_shm_bcast(void* h, void* buf, uint64_t bytes, int root);
int mtb_shm_scatter(void* h, const void* send, void* recv, uint64_t bytes_per_rank, int root);  /* mpipy.py:236-241 */
int mtb_shm_gather(void* h, const void* send, void* recv, uint64_t bytes_per_rank, int root);   /* mpipy.py:121-127 */
int mtb_shm_allgather(void* h, const void* send, void* recv, uint64_t bytes_per_rank);
int mtb_shm_allreduce_f32(void* h, float* buf, uint64_t count, int average);                    /* Gather + np.mean, mpipy.py:121-136 */

/* ---- IDX (optionally gzip) dataset reader (convolutional.extract_data / extract_labels,
 *      imported at mpipy.py:12, used at mpipy.py:215-218) */
int mtb_idx_read_images(const char* path, int64_t n, int64_t rows, int64_t cols,
                        float* out_f32 /* n*rows*cols, (x-127.5)/255 */, uint8_t* out_u8 /* optional raw */);
int mtb_idx_read_labels(const char* path, int64_t n, int64_t* out_i64);
/* CPU restatement of the normalise op, multi-threaded */
void mtb_normalize_u8(const uint8_t* in, float* out, int64_t n);

/* ---- masked CRC32C, used by the TF bundle-V2 checkpoint writer */
uint32_t mtb_crc32c(const void* data, uint64_t n, uint32_t init);

#ifdef __cplusplus
}
#endif
#endif  /* MTB200_HOST_H_ */
