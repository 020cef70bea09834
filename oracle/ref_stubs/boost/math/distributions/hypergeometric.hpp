// Boost.Math stand-in: included by the reference, never used on the path.  TEST INFRASTRUCTURE.
#pragma once
