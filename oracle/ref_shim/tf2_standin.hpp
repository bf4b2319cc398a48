// tf2_standin.hpp — stand-ins for the tf2 pieces the reference's nodes use beyond lookupTransform (which tests/mock_ros2
// provides): tf2_sensor_msgs' doTransform of a PointCloud2 and tf2's LinearMath (Vector3, Quaternion, Matrix3x3,
// Transform, fromMsg).  Test infrastructure only, see pcl_standin.hpp.  Restated third-party arithmetic:
//   tf2::doTransform(PointCloud2)   Eigen::Affine3f from the TransformStamped — translation narrowed to float, rotation
//                                   from Eigen::Quaternionf::toRotationMatrix — applied to each point's x, y, z in
//                                   float32 as (m * p) + t, every other byte of the message copied
//   tf2::Matrix3x3(q), getRPY       the tf2 / Bullet formulas in double (setRotation with s = 2 / |q|^2; getEulerYPR)
#pragma once
#include <cmath>
#include <cstring>

#include "mock_ros2.hpp"

namespace tf2 {

class Vector3 {
public:
    Vector3() = default;
    Vector3(double x, double y, double z) : v_{x, y, z} {}
    double x() const { return v_[0]; }
    double y() const { return v_[1]; }
    double z() const { return v_[2]; }
    double dot(const Vector3& o) const { return v_[0] * o.v_[0] + v_[1] * o.v_[1] + v_[2] * o.v_[2]; }
    void setValue(double x, double y, double z) { v_[0] = x; v_[1] = y; v_[2] = z; }
    Vector3 operator+(const Vector3& o) const { return Vector3(v_[0] + o.v_[0], v_[1] + o.v_[1], v_[2] + o.v_[2]); }
private:
    double v_[3] = {0, 0, 0};
};

class Quaternion {
public:
    Quaternion() = default;
    Quaternion(double x, double y, double z, double w) : q_{x, y, z, w} {}
    double x() const { return q_[0]; }
    double y() const { return q_[1]; }
    double z() const { return q_[2]; }
    double w() const { return q_[3]; }
    double length2() const { return q_[0] * q_[0] + q_[1] * q_[1] + q_[2] * q_[2] + q_[3] * q_[3]; }
    Quaternion inverse() const { return Quaternion(-q_[0], -q_[1], -q_[2], q_[3]); }
    Quaternion operator*(const Quaternion& b) const {              // tf2/LinearMath/Quaternion.h operator*(q1, q2)
        const Quaternion& a = *this;
        return Quaternion(a.w() * b.x() + a.x() * b.w() + a.y() * b.z() - a.z() * b.y(),
                          a.w() * b.y() + a.y() * b.w() + a.z() * b.x() - a.x() * b.z(),
                          a.w() * b.z() + a.z() * b.w() + a.x() * b.y() - a.y() * b.x(),
                          a.w() * b.w() - a.x() * b.x() - a.y() * b.y() - a.z() * b.z());
    }
    void setRPY(double roll, double pitch, double yaw) {          // tf2/LinearMath/Quaternion.h
        const double hy = yaw * 0.5, hp = pitch * 0.5, hr = roll * 0.5;
        const double cy = std::cos(hy), sy = std::sin(hy), cp = std::cos(hp), sp = std::sin(hp), cr = std::cos(hr), sr = std::sin(hr);
        q_[0] = sr * cp * cy - cr * sp * sy;
        q_[1] = cr * sp * cy + sr * cp * sy;
        q_[2] = cr * cp * sy - sr * sp * cy;
        q_[3] = cr * cp * cy + sr * sp * sy;
    }
private:
    double q_[4] = {0, 0, 0, 1};
};

inline Vector3 quatRotate(const Quaternion& q, const Vector3& v) {  // q * v * q^-1, as tf2 (quaternion x pure vector)
    const Quaternion t(q.w() * v.x() + q.y() * v.z() - q.z() * v.y(), q.w() * v.y() + q.z() * v.x() - q.x() * v.z(),
                       q.w() * v.z() + q.x() * v.y() - q.y() * v.x(), -q.x() * v.x() - q.y() * v.y() - q.z() * v.z());
    const Quaternion r = t * q.inverse();
    return Vector3(r.x(), r.y(), r.z());
}

class Matrix3x3 {
public:
    Matrix3x3() = default;
    explicit Matrix3x3(const Quaternion& q) { setRotation(q); }
    void setRotation(const Quaternion& q) {                        // tf2/LinearMath/Matrix3x3.h
        const double d = q.length2(), s = 2.0 / d;
        const double xs = q.x() * s, ys = q.y() * s, zs = q.z() * s;
        const double wx = q.w() * xs, wy = q.w() * ys, wz = q.w() * zs;
        const double xx = q.x() * xs, xy = q.x() * ys, xz = q.x() * zs;
        const double yy = q.y() * ys, yz = q.y() * zs, zz = q.z() * zs;
        m_[0] = Vector3(1.0 - (yy + zz), xy - wz, xz + wy);
        m_[1] = Vector3(xy + wz, 1.0 - (xx + zz), yz - wx);
        m_[2] = Vector3(xz - wy, yz + wx, 1.0 - (xx + yy));
    }
    const Vector3& operator[](int i) const { return m_[i]; }
    void getRPY(double& roll, double& pitch, double& yaw) const {  // getEulerYPR, solution 1
        if (std::fabs(m_[2].x()) >= 1) {
            yaw = 0;
            const double delta = std::atan2(m_[2].y(), m_[2].z());
            if (m_[2].x() < 0) { pitch = M_PI / 2.0; roll = delta; }
            else { pitch = -M_PI / 2.0; roll = delta; }
        } else {
            pitch = -std::asin(m_[2].x());
            roll = std::atan2(m_[2].y() / std::cos(pitch), m_[2].z() / std::cos(pitch));
            yaw = std::atan2(m_[1].x() / std::cos(pitch), m_[0].x() / std::cos(pitch));
        }
    }
private:
    Vector3 m_[3] = {Vector3(1, 0, 0), Vector3(0, 1, 0), Vector3(0, 0, 1)};
};

class Transform {
public:
    Transform() = default;
    void setRotation(const Quaternion& q) { q_ = q; basis_.setRotation(q); }
    void setOrigin(const Vector3& o) { origin_ = o; }
    Quaternion getRotation() const { return q_; }      // the reference only feeds it back into Matrix3x3(q)
    const Vector3& getOrigin() const { return origin_; }
    Vector3 operator*(const Vector3& x) const {
        return Vector3(basis_[0].dot(x) + origin_.x(), basis_[1].dot(x) + origin_.y(), basis_[2].dot(x) + origin_.z());
    }
private:
    Quaternion q_;
    Matrix3x3 basis_;
    Vector3 origin_;
};

inline void fromMsg(const geometry_msgs::msg::Transform& in, Transform& out) {
    out.setOrigin(Vector3(in.translation.x, in.translation.y, in.translation.z));
    out.setRotation(Quaternion(in.rotation.x, in.rotation.y, in.rotation.z, in.rotation.w));
}
inline geometry_msgs::msg::Quaternion toMsg(const Quaternion& q) {
    geometry_msgs::msg::Quaternion m;
    m.x = q.x(); m.y = q.y(); m.z = q.z(); m.w = q.w();
    return m;
}

// tf2_sensor_msgs::doTransform for sensor_msgs::msg::PointCloud2
inline void doTransform(const sensor_msgs::msg::PointCloud2& in, sensor_msgs::msg::PointCloud2& out,
                        const geometry_msgs::msg::TransformStamped& t_in) {
    out = in;
    out.header = t_in.header;
    // Eigen::Quaternionf::toRotationMatrix
    const float qx = static_cast<float>(t_in.transform.rotation.x), qy = static_cast<float>(t_in.transform.rotation.y),
                qz = static_cast<float>(t_in.transform.rotation.z), qw = static_cast<float>(t_in.transform.rotation.w);
    const float tx = 2.0f * qx, ty = 2.0f * qy, tz = 2.0f * qz;
    const float twx = tx * qw, twy = ty * qw, twz = tz * qw;
    const float txx = tx * qx, txy = ty * qx, txz = tz * qx;
    const float tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
    const float m[3][3] = {{1.0f - (tyy + tzz), txy - twz, txz + twy},
                           {txy + twz, 1.0f - (txx + tzz), tyz - twx},
                           {txz - twy, tyz + twx, 1.0f - (txx + tyy)}};
    const float t[3] = {static_cast<float>(t_in.transform.translation.x), static_cast<float>(t_in.transform.translation.y),
                        static_cast<float>(t_in.transform.translation.z)};
    int off[3] = {-1, -1, -1};
    for (const auto& f : in.fields) {
        if (f.name == "x") off[0] = static_cast<int>(f.offset);
        else if (f.name == "y") off[1] = static_cast<int>(f.offset);
        else if (f.name == "z") off[2] = static_cast<int>(f.offset);
    }
    if (off[0] < 0 || off[1] < 0 || off[2] < 0) return;
    const size_t n = static_cast<size_t>(in.width) * in.height;
    for (size_t i = 0; i < n; ++i) {
        float p[3], r[3];
        for (int a = 0; a < 3; ++a) std::memcpy(&p[a], in.data.data() + i * in.point_step + off[a], 4);
        for (int k = 0; k < 3; ++k) {
            float acc = m[k][0] * p[0];
            acc = acc + m[k][1] * p[1];
            acc = acc + m[k][2] * p[2];
            r[k] = acc + t[k];
        }
        for (int a = 0; a < 3; ++a) std::memcpy(out.data.data() + i * in.point_step + off[a], &r[a], 4);
    }
}

}  // namespace tf2
