// empty on purpose: match/flow.cpp includes <xtensor/xsort.hpp>; oracle/ref/flow_driver.cpp supplies the few xt names it uses
