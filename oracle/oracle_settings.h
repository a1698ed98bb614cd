// TEST INFRASTRUCTURE ONLY.  The reference's run-time settings that the restated functions read (util/settings.cpp; the
// reference keeps them as global variables).  One process-wide block, same field order as dso_b200_settings; the macros
// below keep the restatements reading `setting_xxx` like the reference does.
#pragma once
struct dso_oracle_settings
{
	float huberTH;                      // settings.cpp:127
	float outlierTHSumComponent;        // :91
	float maxPixSearch;                 // :156
	int minTraceTestRadius;             // :158
	int GNItsOnPointActivation;         // :159
	float trace_stepsize;               // :160
	int trace_GNIterations;             // :161
	float trace_GNThreshold;            // :162
	float trace_extraSlackOnTH;         // :163
	float trace_slackInterval;          // :164
	float trace_minImprovementFactor;   // :165
	float minIdepthH_act;               // :67
};
extern dso_oracle_settings dso_oracle_set;      // defined in oracle_tracker.cpp
#define setting_huberTH (dso_oracle_set.huberTH)
#define setting_outlierTHSumComponent (dso_oracle_set.outlierTHSumComponent)
#define setting_maxPixSearch (dso_oracle_set.maxPixSearch)
#define setting_minTraceTestRadius (dso_oracle_set.minTraceTestRadius)
#define setting_GNItsOnPointActivation (dso_oracle_set.GNItsOnPointActivation)
#define setting_trace_stepsize (dso_oracle_set.trace_stepsize)
#define setting_trace_GNIterations (dso_oracle_set.trace_GNIterations)
#define setting_trace_GNThreshold (dso_oracle_set.trace_GNThreshold)
#define setting_trace_extraSlackOnTH (dso_oracle_set.trace_extraSlackOnTH)
#define setting_trace_slackInterval (dso_oracle_set.trace_slackInterval)
#define setting_trace_minImprovementFactor (dso_oracle_set.trace_minImprovementFactor)
#define setting_minIdepthH_act (dso_oracle_set.minIdepthH_act)
