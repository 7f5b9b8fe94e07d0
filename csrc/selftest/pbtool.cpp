// psb_pbtool — codec test helper: `psb_pbtool <package.Message>` reads one serialized message
// on stdin, decodes it with the hand-written codec (csrc/wire/pb.h) and writes the re-encoded
// bytes to stdout.  Exit code 3 = decode failure, 2 = unknown message.  tests/test_wire.py
// drives it against the protobuf runtime (golden bytes, unpacked repeated fields, unknown
// fields, negative int32).
#include <iostream>
#include <iterator>
#include <map>
#include <string>

#include "../wire/pb.h"
#include "../wire/pb_ext.h"

template <typename M>
static int roundtrip(const std::string& in) {
  M m;
  if (!m.decode(in)) return 3;
  std::string out = m.encode();
  std::cout.write(out.data(), static_cast<std::streamsize>(out.size()));
  return 0;
}

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  std::string in((std::istreambuf_iterator<char>(std::cin)), std::istreambuf_iterator<char>());
  using namespace psb::pb;
  const std::map<std::string, int (*)(const std::string&)> table = {
      {"coordinator.WorkerInfo", roundtrip<WorkerInfo>},
      {"coordinator.RegisterResponse", roundtrip<RegisterResponse>},
      {"coordinator.HeartbeatRequest", roundtrip<HeartbeatRequest>},
      {"coordinator.HeartbeatResponse", roundtrip<HeartbeatResponse>},
      {"coordinator.ListWorkersRequest", roundtrip<ListWorkersRequest>},
      {"coordinator.ListWorkersResponse", roundtrip<ListWorkersResponse>},
      {"coordinator.GetPSAddressRequest", roundtrip<GetPSAddressRequest>},
      {"coordinator.GetPSAddressResponse", roundtrip<GetPSAddressResponse>},
      {"parameter_server.Tensor", roundtrip<Tensor>},
      {"parameter_server.GradientUpdate", roundtrip<GradientUpdate>},
      {"parameter_server.PushResponse", roundtrip<PushResponse>},
      {"parameter_server.PullRequest", roundtrip<PullRequest>},
      {"parameter_server.ParameterUpdate", roundtrip<ParameterUpdate>},
      {"parameter_server.SyncStatusRequest", roundtrip<SyncStatusRequest>},
      {"parameter_server.SyncStatusResponse", roundtrip<SyncStatusResponse>},
      {"parameter_server.SaveCheckpointRequest", roundtrip<SaveCheckpointRequest>},
      {"parameter_server.SaveCheckpointResponse", roundtrip<SaveCheckpointResponse>},
      {"parameter_server.LoadCheckpointRequest", roundtrip<LoadCheckpointRequest>},
      {"parameter_server.LoadCheckpointResponse", roundtrip<LoadCheckpointResponse>},
      {"psb.ShardInfo", roundtrip<ShardInfo>},
      {"psb.ShardMapResponse", roundtrip<ShardMapResponse>},
      {"psb.BarrierRequest", roundtrip<BarrierRequest>},
      {"psb.BarrierResponse", roundtrip<BarrierResponse>},
      {"psb.ClockRequest", roundtrip<ClockRequest>},
      {"psb.ClockResponse", roundtrip<ClockResponse>},
      {"psb.HandleBlob", roundtrip<HandleBlob>},
      {"psb.HandleQuery", roundtrip<HandleQuery>},
      {"psb.HandleList", roundtrip<HandleList>},
      {"psb.Ack", roundtrip<Ack>},
      {"psb.TableSpec", roundtrip<TableSpec>},
      {"psb.SparseRows", roundtrip<SparseRows>},
      {"psb.SparsePush", roundtrip<SparsePush>},
      {"psb.SparsePushResponse", roundtrip<SparsePushResponse>},
      {"psb.RowsRequest", roundtrip<RowsRequest>},
      {"psb.RowsResponse", roundtrip<RowsResponse>},
  };
  auto it = table.find(argv[1]);
  if (it == table.end()) return 2;
  return it->second(in);
}
