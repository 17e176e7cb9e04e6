/* b2d_hostcopy.h — multi-threaded host memcpy with non-temporal stores (b2d_hostcopy.cpp). Not installed. */
#ifndef B2D_HOSTCOPY_H
#define B2D_HOSTCOPY_H
#include <stddef.h>
/* copy n bytes with up to nt threads (the caller's included); small copies run on the calling thread alone */
void b2d_par_memcpy(void *dst, const void *src, size_t n, unsigned nt);
#endif
