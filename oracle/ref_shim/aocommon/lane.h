// ORACLE — TEST INFRASTRUCTURE ONLY. aocommon/lane.h is included by the reference but no
// Lane is used on the compiled path; intentionally empty (ours).
#ifndef AOCOMMON_LANE_H_
#define AOCOMMON_LANE_H_
#endif
