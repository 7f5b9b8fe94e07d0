/* The native coordinator client from C (include/psb_control.h, libpsb_control.so): what the
 * Python package does through ctypes, and what a C / C++ / Fortran trainer would call directly.
 * It registers as a worker (the reference's bootstrap, /root/reference/src/worker.cpp:108-186),
 * announces a shard, meets its peers at a barrier, exchanges a small blob with them and reads the
 * registry.
 *
 *   build/bin/coordinator 127.0.0.1:50052 127.0.0.1:50051 &
 *   gcc -std=c11 -I include examples/control_client.c \
 *       -L parameter_server_distributed_b200/_native -lpsb_control \
 *       -Wl,-rpath,$PWD/parameter_server_distributed_b200/_native -o /tmp/control_client
 *   /tmp/control_client 127.0.0.1:50052 0 2 &  /tmp/control_client 127.0.0.1:50052 1 2
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "psb_control.h"

int main(int argc, char** argv) {
  const char* addr = argc > 1 ? argv[1] : "127.0.0.1:50052";
  const int rank = argc > 2 ? atoi(argv[2]) : 0;
  const int world = argc > 3 ? atoi(argv[3]) : 1;

  PsbCoord* c = psbc_connect(addr, rank, "127.0.0.1", 0, "c-client", /*heartbeat_ms=*/200,
                             /*connect_timeout_ms=*/10000);
  if (!c) {
    fprintf(stderr, "connect failed: %s\n", psbc_last_error());
    return 1;
  }
  printf("rank %d: registered; parameter server at %s\n", rank, psbc_ps_address(c));

  char shard_addr[64];
  snprintf(shard_addr, sizeof shard_addr, "hbm://example/gpu%d", rank);
  if (psbc_register_shard(c, rank, shard_addr, rank) != 0) goto fail;
  if (psbc_barrier(c, "example/up", world, 10000) != 0) goto fail;

  /* every rank contributes 16 bytes, every rank receives world x 16 bytes */
  char mine[16], all[16 * 64];
  memset(mine, 0, sizeof mine);
  snprintf(mine, sizeof mine, "hello from %d", rank);
  if (world > 64 || psbc_allgather(c, "example/greeting", mine, sizeof mine, world, all, 10000) != 0)
    goto fail;
  for (int r = 0; r < world; ++r) printf("rank %d: rank %d says '%s'\n", rank, r, all + 16 * r);

  int32_t ids[64], count = 0;
  if (psbc_list_workers(c, ids, 64, &count) != 0) goto fail;
  printf("rank %d: %d worker(s) registered\n", rank, count);

  int32_t slowest = -1;
  if (psbc_update_clock(c, 5 + rank, &slowest) != 0) goto fail;
  printf("rank %d: my clock %d, slowest live clock %d\n", rank, 5 + rank, slowest);

  if (psbc_barrier(c, "example/down", world, 10000) != 0) goto fail;
  psbc_close(c);
  return 0;
fail:
  fprintf(stderr, "rank %d: %s\n", rank, psbc_last_error());
  psbc_close(c);
  return 1;
}
