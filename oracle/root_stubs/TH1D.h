#include <TH1.h>
