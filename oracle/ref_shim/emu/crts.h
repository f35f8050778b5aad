/* Sunway <crts.h>: nothing from it is used by the stencil kernels (see athread.h). */
