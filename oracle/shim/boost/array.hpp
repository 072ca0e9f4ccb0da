// Shim for <boost/array.hpp>: included by ccicp2d/kdtree.h:16 but unused there.
#pragma once
