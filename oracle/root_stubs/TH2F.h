// empty stand-in: nothing from <TH2F.h> is used on the hot path
