// oracle/stubs/geometry/point.h -- TEST INFRASTRUCTURE ONLY.
// Stand-in for FLIRTLib's <geometry/point.h>, which the reference's src/mapGraph/PMap.h includes and which is not in
// the reference tree.  PMap.cpp / line.cpp only need a 2-D point with double coordinates and the standard headers
// FLIRTLib's header pulls in; nothing here is taken from the reference.
#ifndef ORC_STUB_GEOMETRY_POINT_H
#define ORC_STUB_GEOMETRY_POINT_H
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <iostream>
struct Point2D {
    double x, y;
    Point2D() : x(0.0), y(0.0) {}
    Point2D(double _x, double _y) : x(_x), y(_y) {}
};
#endif
