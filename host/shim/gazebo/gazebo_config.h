// empty: opende's quickstep_util.h includes it; nothing from it is used
