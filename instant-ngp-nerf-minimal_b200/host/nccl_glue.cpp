// NCCL communicator bootstrap used by Testbed::set_data_parallel.
#include "testbed.h"
#include <nccl.h>

std::string ngp_host_nccl_unique_id() {
	ncclUniqueId id;
	const ncclResult_t r = ncclGetUniqueId(&id);
	if (r != ncclSuccess) throw std::runtime_error(std::string("ncclGetUniqueId failed: ") + ncclGetErrorString(r));
	return std::string((const char*)&id, sizeof(id));
}

void* ngp_host_nccl_comm_init(const std::string& id_bytes, int rank, int world_size) {
	if (id_bytes.size() != sizeof(ncclUniqueId)) throw std::runtime_error("nccl_comm_init: bad unique id size");
	ncclUniqueId id;
	memcpy(&id, id_bytes.data(), sizeof(id));
	ncclComm_t comm = nullptr;
	const ncclResult_t r = ncclCommInitRank(&comm, world_size, id, rank);
	if (r != ncclSuccess) throw std::runtime_error(std::string("ncclCommInitRank failed: ") + ncclGetErrorString(r));
	return comm;
}

void ngp_host_nccl_comm_destroy(void* comm) { if (comm) ncclCommDestroy((ncclComm_t)comm); }
