/* Emulator of the subset of the Sunway vendor header <athread.h> the reference's stencil
 * kernels use (src/stencil/slave/*.cpp), so that the reference's OWN accelerator kernels —
 * not only its naive check loop — can be executed here, unmodified, from where they lie under
 * /root/reference.  An 8x8 array of "CPEs" is 64 pthreads; each owns an arena that holds its
 * stack and its ldm_malloc heap, so a local address (stack variable or LDM block) names the
 * same object in every CPE's arena — which is what athread_rma_iput's remote addresses mean
 * on the real machine (stencil_rma.cpp:198-255).  DMA is synchronous (memcpy + reply
 * increment), athread_ssync_array is a barrier.  Semantics restated from the call sites:
 *   *_stride(dst, src, len, bsize, stride): `len` bytes in runs of `bsize`; the MAIN-MEMORY
 *   side skips `stride` bytes between runs, the LDM side is dense (stencil_dma.cpp:126-150).
 * Test infrastructure only (oracle/): never linked into the product library. */
#ifndef SB200_REF_EMU_ATHREAD_H
#define SB200_REF_EMU_ATHREAD_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef volatile int athread_rply_t;

/* ---- host side (src/stencil/stencil.cpp:33-54, src/main.cpp:54) ---- */
#ifndef __sw_slave__
#define SLAVE_FUN(name) slave_##name
#endif
int athread_init(void);
int athread_spawn(void* fn, void* arg);
int athread_join(void);
int sb200_emu_last_spawn_failed(void); /* emulator only: the CPE stacks were not congruent */

/* ---- CPE side ---- */
extern __thread int sb200_emu_row, sb200_emu_col;
#define _ROW sb200_emu_row
#define _COL sb200_emu_col

void* ldm_malloc(size_t bytes);
void ldm_free(void* ptr, size_t bytes);

void athread_dma_get(void* dst_ldm, void* src_mem, size_t len);
void athread_dma_put(void* dst_mem, void* src_ldm, size_t len);
void athread_dma_get_stride(void* dst_ldm, void* src_mem, size_t len, size_t bsize, size_t stride);
void athread_dma_put_stride(void* dst_mem, void* src_ldm, size_t len, size_t bsize, size_t stride);
void athread_dma_iget(void* dst_ldm, void* src_mem, size_t len, athread_rply_t* reply);
void athread_dma_iput(void* dst_mem, void* src_ldm, size_t len, athread_rply_t* reply);
void athread_dma_iget_stride(void* dst_ldm, void* src_mem, size_t len, size_t bsize, size_t stride,
                             athread_rply_t* reply);
void athread_dma_iput_stride(void* dst_mem, void* src_ldm, size_t len, size_t bsize, size_t stride,
                             athread_rply_t* reply);
void athread_dma_wait_value(athread_rply_t* reply, int value);

void athread_rma_iput(void* src_ldm, athread_rply_t* local_reply, size_t len, int remote_id,
                      void* remote_dst_ldm, athread_rply_t* remote_reply);
void athread_rma_wait_value(athread_rply_t* reply, int value);

void athread_ssync_array(void);

#ifdef __cplusplus
}
#endif
#endif
