// tracker_host.cpp -- host half of the ICP depth tracker: the Levenberg-Marquardt loop of
// ITMDepthTracker::TrackCamera (ITMLib/Engine/ITMDepthTracker.cpp:153-211) around the device sums of tracker.cu.
//
// A maintainer who subclasses ITMDepthTracker keeps the reference's own loop and needs only
// dsb200_tracker_compute_g_and_h; this file is for callers without the reference (the Python mirror, bench.py),
// so that "track a frame" is one C call.  Everything here is a few dozen flops per iteration in float, compiled
// by g++ with -ffp-contract=off like hostmath.cpp so that it follows the reference's rounding:
//   pose <-> 6 parameters   ITMLib/Objects/ITMPose.cpp:82-262 (SetModelViewFromParams / SetParamsFromModelView)
//   Cholesky solve          ORUtils/Cholesky.h:11-72
//   step -> increment       ITMDepthTracker.cpp:122-151 (ApplyDelta), :111-120 (HasConverged)
#include "../../include/dsb200.h"
#include <math.h>
#include <string.h>

namespace dsb { int fail_with(int code, const char *msg); }   // capi.cu: records the text dsb200_last_error() returns

namespace {

inline float dot3(const float *a, const float *b)
{
    float r = 0;
    for (int i = 0; i < 3; i++) r += a[i] * b[i];
    return r;
}
inline void cross3(const float *a, const float *b, float *r)
{
    r[0] = a[1] * b[2] - a[2] * b[1];
    r[1] = a[2] * b[0] - a[0] * b[2];
    r[2] = a[0] * b[1] - a[1] * b[0];
}

// params = (tx, ty, tz, rx, ry, rz) -> column-major 4x4 (exponential map, three ranges of the rotation angle)
void model_view_from_params(const float *p, float *M)
{
    const float one_6th = 1.0f / 6.0f, one_20th = 1.0f / 20.0f;
    const float *t = p, *w = p + 3;
    const float theta_sq = dot3(w, w);
    const float theta = sqrtf(theta_sq);
    float A, B, T[3], wxt[3];
    cross3(w, t, wxt);
    if (theta_sq < 1e-8f) {
        A = 1.0f - one_6th * theta_sq; B = 0.5f;
        for (int i = 0; i < 3; i++) T[i] = t[i] + 0.5f * wxt[i];
    } else {
        float C;
        if (theta_sq < 1e-6f) {
            C = one_6th * (1.0f - one_20th * theta_sq);
            A = 1.0f - theta_sq * C;
            B = 0.5f - 0.25f * one_6th * theta_sq;
        } else {
            const float inv_theta = 1.0f / theta;
            A = sinf(theta) * inv_theta;
            B = (1.0f - cosf(theta)) * (inv_theta * inv_theta);
            C = (1.0f - A) * (inv_theta * inv_theta);
        }
        float wxwxt[3];
        cross3(w, wxt, wxwxt);
        for (int i = 0; i < 3; i++) T[i] = t[i] + B * wxt[i] + C * wxwxt[i];
    }
    const float wx2 = w[0] * w[0], wy2 = w[1] * w[1], wz2 = w[2] * w[2];
    float R[9];   // R[row + 3 * col]
    R[0 + 3 * 0] = 1.0f - B * (wy2 + wz2);
    R[1 + 3 * 1] = 1.0f - B * (wx2 + wz2);
    R[2 + 3 * 2] = 1.0f - B * (wx2 + wy2);
    float a, b;
    a = A * w[2]; b = B * (w[0] * w[1]);
    R[0 + 3 * 1] = b - a; R[1 + 3 * 0] = b + a;
    a = A * w[1]; b = B * (w[0] * w[2]);
    R[0 + 3 * 2] = b + a; R[2 + 3 * 0] = b - a;
    a = A * w[0]; b = B * (w[1] * w[2]);
    R[1 + 3 * 2] = b - a; R[2 + 3 * 1] = b + a;
    for (int c = 0; c < 3; c++) {
        for (int r = 0; r < 3; r++) M[r + 4 * c] = R[r + 3 * c];
        M[3 + 4 * c] = 0.0f;
    }
    M[12] = T[0]; M[13] = T[1]; M[14] = T[2]; M[15] = 1.0f;
}

// column-major 4x4 -> params (logarithm; the near-pi branch picks the rotation axis from the symmetric part)
void params_from_model_view(const float *M, float *p)
{
    float R[9], T[3] = {M[12], M[13], M[14]};
    for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) R[r + 3 * c] = M[r + 4 * c];
    float rot[3];
    const float cos_angle = (R[0] + R[4] + R[8] - 1.0f) * 0.5f;
    rot[0] = (R[2 + 3 * 1] - R[1 + 3 * 2]) * 0.5f;
    rot[1] = (R[0 + 3 * 2] - R[2 + 3 * 0]) * 0.5f;
    rot[2] = (R[1 + 3 * 0] - R[0 + 3 * 1]) * 0.5f;
    const float sin_angle_abs = sqrtf(dot3(rot, rot));
    if (cos_angle > M_SQRT1_2) {
        if (sin_angle_abs) {
            const float s = asinf(sin_angle_abs) / sin_angle_abs;
            for (int i = 0; i < 3; i++) rot[i] *= s;
        }
    } else if (cos_angle > -M_SQRT1_2) {
        const float s = acosf(cos_angle) / sin_angle_abs;
        for (int i = 0; i < 3; i++) rot[i] *= s;
    } else {
        const float angle = (float)M_PI - asinf(sin_angle_abs);
        const float d0 = R[0] - cos_angle, d1 = R[4] - cos_angle, d2 = R[8] - cos_angle;
        float r2[3];
        if (fabsf(d0) > fabsf(d1) && fabsf(d0) > fabsf(d2)) {
            r2[0] = d0; r2[1] = (R[1 + 3 * 0] + R[0 + 3 * 1]) * 0.5f; r2[2] = (R[0 + 3 * 2] + R[2 + 3 * 0]) * 0.5f;
        } else if (fabsf(d1) > fabsf(d2)) {
            r2[0] = (R[1 + 3 * 0] + R[0 + 3 * 1]) * 0.5f; r2[1] = d1; r2[2] = (R[2 + 3 * 1] + R[1 + 3 * 2]) * 0.5f;
        } else {
            r2[0] = (R[0 + 3 * 2] + R[2 + 3 * 0]) * 0.5f; r2[1] = (R[2 + 3 * 1] + R[1 + 3 * 2]) * 0.5f; r2[2] = d2;
        }
        if (dot3(r2, rot) < 0.0f) for (int i = 0; i < 3; i++) r2[i] *= -1.0f;
        const float len = sqrtf(dot3(r2, r2));
        for (int i = 0; i < 3; i++) r2[i] = len == 0 ? 0.0f : r2[i] / len;
        for (int i = 0; i < 3; i++) rot[i] = angle * r2[i];
    }

    float shtot = 0.5f;
    const float theta = sqrtf(dot3(rot, rot));
    if (theta > 0.00001f) shtot = sinf(theta * 0.5f) / theta;

    // rotate the translation back by half the angle
    float half[6] = {0.0f, 0.0f, 0.0f, rot[0] * -0.5f, rot[1] * -0.5f, rot[2] * -0.5f}, H[16];
    model_view_from_params(half, H);
    float rt[3];
    for (int r = 0; r < 3; r++) rt[r] = H[r + 4 * 0] * T[0] + H[r + 4 * 1] * T[1] + H[r + 4 * 2] * T[2];

    if (theta > 0.001f) {
        const float denom = dot3(rot, rot);
        const float param = dot3(T, rot) * (1 - 2 * shtot) / denom;
        for (int i = 0; i < 3; i++) rt[i] -= rot[i] * param;
    } else {
        const float param = dot3(T, rot) / 24;
        for (int i = 0; i < 3; i++) rt[i] -= rot[i] * param;
    }
    for (int i = 0; i < 3; i++) rt[i] /= 2 * shtot;
    p[0] = rt[0]; p[1] = rt[1]; p[2] = rt[2]; p[3] = rot[0]; p[4] = rot[1]; p[5] = rot[2];
}

// ORUtils::Cholesky<float>(mat, n) followed by Backsub(x, v): in-place LDL-style factor, forward, scale, backward
void cholesky_solve(const float *mat, int n, const float *v, float *x)
{
    float L[36], y[6];
    for (int i = 0; i < n * n; i++) L[i] = mat[i];
    for (int c = 0; c < n; c++) {
        float inv_diag = 1;
        for (int r = c; r < n; r++) {
            float val = L[c + r * n];
            for (int c2 = 0; c2 < c; c2++) val -= L[c + c2 * n] * L[c2 + r * n];
            if (r == c) { L[c + r * n] = val; inv_diag = 1.0f / val; }
            else { L[r + c * n] = val; L[c + r * n] = val * inv_diag; }
        }
    }
    for (int i = 0; i < n; i++) {
        float val = v[i];
        for (int j = 0; j < i; j++) val -= L[j + i * n] * y[j];
        y[i] = val;
    }
    for (int i = 0; i < n; i++) y[i] /= L[i + i * n];
    for (int i = n - 1; i >= 0; i--) {
        float val = y[i];
        for (int j = i + 1; j < n; j++) val -= L[i + j * n] * x[j];
        x[i] = val;
    }
}

} // namespace

extern "C" {

int dsb200_pose_coerce(float *M, float *params)
{
    if (!M) return dsb::fail_with(DSB200_ERR_INVALID, "pose_coerce: null argument");
    float p[6];
    params_from_model_view(M, p);
    model_view_from_params(p, M);
    if (params) memcpy(params, p, sizeof(p));
    return DSB200_OK;
}

int dsb200_pose_from_params(const float *params, float *M)
{
    if (!params || !M) return dsb::fail_with(DSB200_ERR_INVALID, "pose_from_params: null argument");
    model_view_from_params(params, M);
    return DSB200_OK;
}

int dsb200_tracker_track_camera(dsb200_tracker *t, const dsb200_tracker_config *cfg, const dsb200_view *view,
                                dsb200_tracking *trk, const float *scenePose, int32_t sceneW, int32_t sceneH,
                                dsb200_tracker_report *report)
{
    if (!t || !cfg || !view || !trk || !scenePose || !view->depth || !trk->icp_points || !trk->icp_normals)
        return dsb::fail_with(DSB200_ERR_INVALID, "tracker_track_camera: null argument");
    if (cfg->levels < 1 || cfg->levels > DSB200_TRACKER_MAX_LEVELS || cfg->no_icp_run_till_level < 0)
        return dsb::fail_with(DSB200_ERR_INVALID, "tracker_track_camera: bad level count");
    // SetEvaluationData + PrepareForEvaluation (:43-76)
    int rc = dsb200_tracker_set_evaluation_data(t, view->depth, view->intr_d, trk->icp_points, trk->icp_normals, sceneW, sceneH,
                                                view->intr_d, scenePose);
    if (rc) return rc;

    // distance thresholds per level, ITMDepthTracker.cpp:26-30; iterations per level :20-24
    float distThresh[DSB200_TRACKER_MAX_LEVELS]; int noIterations[DSB200_TRACKER_MAX_LEVELS];
    const float step = cfg->dist_thresh / cfg->levels;
    distThresh[cfg->levels - 1] = cfg->dist_thresh;
    for (int l = cfg->levels - 2; l >= 0; l--) distThresh[l] = distThresh[l + 1] - step;
    noIterations[0] = 2;
    for (int l = 1; l < cfg->levels; l++) noIterations[l] = noIterations[l - 1] + 2;

    float M[16], hessianGood[36], nablaGood[6];
    memcpy(M, trk->M_d, sizeof(M));
    memset(hessianGood, 0, sizeof(hessianGood)); memset(nablaGood, 0, sizeof(nablaGood));
    bool haveGood = false;     // the reference reads hessian_good uninitialised if the very first evaluation fails
    if (report) memset(report, 0, sizeof(*report));

    for (int level = cfg->levels - 1; level >= cfg->no_icp_run_till_level; level--) {
        const int type = cfg->regime[level];
        if (type == DSB200_TRACKER_ITERATION_NONE) continue;
        const bool shortIt = type != DSB200_TRACKER_ITERATION_BOTH;
        float approxInvPose[16], goodM[16];
        if (!dsb200_mat4_inverse(M, approxInvPose)) return dsb::fail_with(DSB200_ERR_INVALID, "tracker_track_camera: singular start pose");
        memcpy(goodM, M, sizeof(M));
        float fOld = 1e20f, lambda = 1.0f;
        for (int it = 0; it < noIterations[level]; it++) {
            float fNew = 0.0f, nablaNew[6] = {0}, hessianNew[36] = {0}; int32_t valid = 0;
            rc = dsb200_tracker_compute_g_and_h(t, level, type, distThresh[level], approxInvPose, &fNew, nablaNew, hessianNew, &valid);
            if (rc) return rc;
            if (report) { report->evaluations++; report->last_valid_points = valid; report->last_f = fNew; }
            if (valid <= 0 || fNew > fOld) {
                memcpy(M, goodM, sizeof(M));
                dsb200_mat4_inverse(M, approxInvPose);
                lambda *= 10.0f;
            } else {
                memcpy(goodM, M, sizeof(M));
                fOld = fNew;
                for (int i = 0; i < 36; i++) hessianGood[i] = hessianNew[i] / valid;
                for (int i = 0; i < 6; i++) nablaGood[i] = nablaNew[i] / valid;
                lambda /= 10.0f;
                haveGood = true;
            }
            if (!haveGood) break;      // nothing to step along yet: leave the pose alone at this level
            float A[36], delta[6] = {0, 0, 0, 0, 0, 0};
            for (int i = 0; i < 36; i++) A[i] = hessianGood[i];
            for (int i = 0; i < 6; i++) A[i + i * 6] *= 1.0f + lambda;
            // ComputeDelta :86-109
            if (shortIt) {
                float small[9];
                for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) small[r + c * 3] = A[r + c * 6];
                cholesky_solve(small, 3, nablaGood, delta);
            } else cholesky_solve(A, 6, nablaGood, delta);
            // ApplyDelta :122-151
            float s[6] = {0, 0, 0, 0, 0, 0};
            if (type == DSB200_TRACKER_ITERATION_ROTATION) { s[0] = delta[0]; s[1] = delta[1]; s[2] = delta[2]; }
            else if (type == DSB200_TRACKER_ITERATION_TRANSLATION) { s[3] = delta[0]; s[4] = delta[1]; s[5] = delta[2]; }
            else for (int i = 0; i < 6; i++) s[i] = delta[i];
            float Tinc[16] = {1.0f, -s[2], s[1], 0.0f, s[2], 1.0f, -s[0], 0.0f, -s[1], s[0], 1.0f, 0.0f, s[3], s[4], s[5], 1.0f};
            float next[16];
            dsb200_mat4_multiply(Tinc, approxInvPose, next);
            // pose_d->SetInvM(next); Coerce(); GetInvM()
            if (!dsb200_mat4_inverse(next, M)) return dsb::fail_with(DSB200_ERR_INVALID, "tracker_track_camera: singular pose update");
            dsb200_pose_coerce(M, nullptr);
            dsb200_mat4_inverse(M, approxInvPose);
            if (report) report->steps++;
            // HasConverged :111-120 (on the solver's step, before the regime scatter)
            float len = 0.0f;
            for (int i = 0; i < 6; i++) len += delta[i] * delta[i];
            if (sqrtf(len) / 6 < cfg->termination_threshold) break;
        }
    }
    memcpy(trk->M_d, M, sizeof(M));
    dsb200_mat4_inverse(M, trk->invM_d);
    return DSB200_OK;
}

} // extern "C"
