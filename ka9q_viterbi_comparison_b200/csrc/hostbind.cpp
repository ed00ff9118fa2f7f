// hostbind.cpp -- place the calling host thread next to a GPU (one host thread / process per device feeds
// its shard over PCIe; the path is PCIe-bound end to end, so the pinned staging buffers must live on the memory node
// the GPU's root complex hangs off, and the thread that issues the copies should run there too).
// Reads the topology from sysfs (no libnuma dependency): /sys/bus/pci/devices/<bdf>/{numa_node,local_cpulist}.
#include "../../include/b200_viterbi.h"

#include <cuda_runtime.h>
#include <sched.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

namespace {

bool read_line(const std::string& path, std::string& out) {
    FILE* f = fopen(path.c_str(), "r");
    if (!f) return false;
    char buf[4096];
    const bool ok = fgets(buf, sizeof(buf), f) != nullptr;
    fclose(f);
    if (!ok) return false;
    out = buf;
    while (!out.empty() && isspace((unsigned char)out.back())) out.pop_back();
    return true;
}

// "0-15,32-47" -> cpu_set_t; returns the number of CPUs set
int parse_cpulist(const std::string& s, cpu_set_t* set) {
    CPU_ZERO(set);
    int n = 0;
    const char* p = s.c_str();
    while (*p) {
        char* e;
        long a = strtol(p, &e, 10);
        if (e == p) break;
        long b = a;
        p = e;
        if (*p == '-') { b = strtol(p + 1, &e, 10); p = e; }
        for (long c = a; c <= b && c < CPU_SETSIZE; c++) { CPU_SET((int)c, set); n++; }
        if (*p == ',') p++;
    }
    return n;
}

}  // namespace

// Returns the NUMA node the calling thread was bound to (>= 0), or -1 when the topology does not say (single node,
// virtualised sysfs, no permission): then nothing is changed. Never fails the caller.
extern "C" int b200v_bind_host_to_device(int device) {
    char bdf[32] = "";
    if (cudaDeviceGetPCIBusId(bdf, (int)sizeof(bdf), device) != cudaSuccess) { cudaGetLastError(); return -1; }
    for (char* c = bdf; *c; c++) *c = (char)tolower((unsigned char)*c);
    const std::string dir = std::string("/sys/bus/pci/devices/") + bdf;
    std::string s;
    if (!read_line(dir + "/numa_node", s)) return -1;
    const int node = atoi(s.c_str());
    if (node < 0) return -1;
    // CPUs of that node that this process may use
    cpu_set_t want, have, both;
    std::string cpus;
    if (!read_line(dir + "/local_cpulist", cpus) || parse_cpulist(cpus, &want) == 0) {
        if (!read_line("/sys/devices/system/node/node" + std::to_string(node) + "/cpulist", cpus) ||
            parse_cpulist(cpus, &want) == 0) return -1;
    }
    if (sched_getaffinity(0, sizeof(have), &have) != 0) return -1;
    CPU_AND(&both, &want, &have);
    if (CPU_COUNT(&both) > 0) sched_setaffinity(0, sizeof(both), &both);
    // prefer the node's memory for everything this thread allocates from now on (cudaMallocHost included);
    // MPOL_PREFERRED = 1. Containers may refuse the call: the CPU binding alone still gives first-touch locality.
    if (node < 1024) {
        unsigned long mask[16] = { 0 };
        mask[node / (8 * sizeof(unsigned long))] |= 1ul << (node % (8 * sizeof(unsigned long)));
#ifdef SYS_set_mempolicy
        (void)syscall(SYS_set_mempolicy, 1, mask, (unsigned long)(sizeof(mask) * 8));
#endif
    }
    return node;
}
