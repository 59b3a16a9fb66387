// empty stand-in: nothing from <TMath.h> is used on the hot path
