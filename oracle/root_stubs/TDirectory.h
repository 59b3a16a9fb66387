#ifndef SXMC_STUB_TDIRECTORY_H
#define SXMC_STUB_TDIRECTORY_H
#include <map>
#include <string>
struct TDirectory {
  std::map<std::string, void*> objects;
  void* Get(const char* name) { return objects.count(name) ? objects[name] : 0; }
};
extern TDirectory* gDirectory;
#endif
