// wrapper TU: reference kpkt.cc compiled unmodified + accessor to its file-static cooling list
#include "kpkt.cc"
#include "refglue.h"
namespace refglue {
void coolinglist_entry(int i, int *type, int *level, int *upperlevel) {
  *type = kpkt::coolinglist[i].type;
  *level = kpkt::coolinglist[i].level;
  *upperlevel = kpkt::coolinglist[i].upperlevel;
}
}  // namespace refglue
