// wrapper TU: reference update_packets.cc compiled unmodified, with its entry point renamed so that the hook
// (hook.cc) can stand at the reference's own call site sn3d.cc:574 -- the drop-in boundary.
#define update_packets ref_update_packets
#include "update_packets.cc"
