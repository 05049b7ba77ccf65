/* Layout of the shared mapping used by the MPI stand-in. */
#ifndef SA_SHIM_COMMON_H_
#define SA_SHIM_COMMON_H_
#include <pthread.h>
#include <stddef.h>

struct ShimControl {
    pthread_barrier_t barrier; /* process-shared, count = world size */
    int world;
    size_t data_bytes;
};
#define SHIM_DATA_OFFSET 4096
#endif
