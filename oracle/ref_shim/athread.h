/* Stub of the Sunway vendor header <athread.h> (absent from this image), just enough
 * to compile the reference's HOST sources unmodified from /root/reference.  The
 * accelerator spawn/join are no-ops: only Stencil::check_result (the naive CPU loop,
 * src/stencil/stencil.cpp:75-151) is exercised.  Test infrastructure only. */
#ifndef SB200_REF_SHIM_ATHREAD_H
#define SB200_REF_SHIM_ATHREAD_H
#define SLAVE_FUN(name) slave_##name
static inline int athread_init(void) { return 0; }
static inline int athread_spawn(void* fn, void* arg) { (void)fn; (void)arg; return 0; }
static inline int athread_join(void) { return 0; }
#endif
