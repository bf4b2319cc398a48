// GeographicLib/LocalCartesian.hpp — stand-in so that the reference's pointcloud_merger.cpp compiles here (GeographicLib
// is not in this image).  Test infrastructure only, see pcl_standin.hpp.  Only the GNSS -> TF half of that node
// (src/pointcloud_merger.cpp:185-305, out of scope) calls it; the merge itself (:308-394) never does,
// so the fixtures made with it (tests/golden/refsrc_merger.npz) do not depend on this arithmetic.  Restated from the
// published definition: WGS84 geodetic -> ECEF, then the east-north-up rotation at the origin, all in double.  Not
// claimed bit-identical to GeographicLib (which evaluates the same map through Geocentric::IntForward).
#pragma once
#include <cmath>

namespace GeographicLib {

class LocalCartesian {
public:
    LocalCartesian() { Reset(0.0, 0.0, 0.0); }
    LocalCartesian(double lat0, double lon0, double h0 = 0.0) { Reset(lat0, lon0, h0); }
    void Reset(double lat0, double lon0, double h0 = 0.0) {
        lat0_ = lat0; lon0_ = lon0; h0_ = h0;
        ecef(lat0, lon0, h0, x0_, y0_, z0_);
        const double d = M_PI / 180.0, sp = std::sin(lat0 * d), cp = std::cos(lat0 * d), sl = std::sin(lon0 * d), cl = std::cos(lon0 * d);
        r_[0][0] = -sl;      r_[0][1] = cl;       r_[0][2] = 0.0;      // east
        r_[1][0] = -sp * cl; r_[1][1] = -sp * sl; r_[1][2] = cp;       // north
        r_[2][0] = cp * cl;  r_[2][1] = cp * sl;  r_[2][2] = sp;       // up
    }
    void Forward(double lat, double lon, double h, double& x, double& y, double& z) const {
        double X, Y, Z;
        ecef(lat, lon, h, X, Y, Z);
        X -= x0_; Y -= y0_; Z -= z0_;
        x = r_[0][0] * X + r_[0][1] * Y + r_[0][2] * Z;
        y = r_[1][0] * X + r_[1][1] * Y + r_[1][2] * Z;
        z = r_[2][0] * X + r_[2][1] * Y + r_[2][2] * Z;
    }
    double LatitudeOrigin() const { return lat0_; }
    double LongitudeOrigin() const { return lon0_; }
    double HeightOrigin() const { return h0_; }

private:
    static void ecef(double lat, double lon, double h, double& X, double& Y, double& Z) {
        const double a = 6378137.0, f = 1.0 / 298.257223563, e2 = f * (2.0 - f), d = M_PI / 180.0;
        const double sp = std::sin(lat * d), cp = std::cos(lat * d), n = a / std::sqrt(1.0 - e2 * sp * sp);
        X = (n + h) * cp * std::cos(lon * d);
        Y = (n + h) * cp * std::sin(lon * d);
        Z = (n * (1.0 - e2) + h) * sp;
    }
    double lat0_ = 0, lon0_ = 0, h0_ = 0, x0_ = 0, y0_ = 0, z0_ = 0, r_[3][3] = {};
};

}  // namespace GeographicLib
