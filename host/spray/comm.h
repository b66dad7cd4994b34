// comm.h — process identity and the NCCL communicator of the host driver.  Replaces the reference's
// `spray::mpi` helpers (src/utils/comm.h: rank(), size(), isRootProcess(), isSingleProcess()) and
// MPI_Init_thread / MPI_Comm_rank (src/apps/spray_ooc.cc:33-37).  One process per GPU on ONE box, launched
// by torchrun-style environment variables (RANK, WORLD_SIZE, LOCAL_RANK, MASTER_PORT); the 128-byte NCCL
// unique id travels through a file in /dev/shm (there is no MPI here to broadcast it).
#pragma once
#include <string>

#include "spray_gpu.h"

namespace spray {
namespace mpi {

void init();  // reads RANK / WORLD_SIZE / LOCAL_RANK (defaults 0 / 1 / 0)
int rank();
int size();
int localRank();
inline bool isRootProcess() { return rank() == 0; }
inline bool isSingleProcess() { return size() == 1; }

// collective over all ranks: creates the NCCL communicator on `dev` (nullptr when size() == 1)
spraygpu_comm createComm(spraygpu_device dev);

}  // namespace mpi
}  // namespace spray
