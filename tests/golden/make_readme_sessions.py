"""Extracts the reference's hand-typed `nc` sessions (the only golden vectors it has, SURVEY.md §4
G1-G8) from /root/reference/chapter*/README.md into tests/golden/readme_sessions.json.

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_readme_sessions.py
The line ranges are the citations used in SURVEY.md; the script fails if a block moved.
"""
import json
import os

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "readme_sessions.json")


def block(path, first, last):
    """Lines first..last (1-based, inclusive) of a README, stripped, blanks and fences dropped."""
    with open(os.path.join(REF, path), encoding="utf-8") as f:
        lines = f.read().split("\n")
    got = [l.strip() for l in lines[first - 1:last]]
    return [l for l in got if l and not l.startswith("```")]


def main():
    g = {}
    g["G1_filter"] = {"cite": "chapter1/README.md:114-123 (predicate Main.java:31)",
                      "input": block("chapter1/README.md", 116, 117), "output": block("chapter1/README.md", 122, 122)}
    g["G2_passthrough"] = {"cite": "chapter1/README.md:72-84",
                           "input": block("chapter1/README.md", 73, 75), "output": block("chapter1/README.md", 81, 83)}
    g["G3_rolling_max"] = {"cite": "chapter2/README.md:54-66 (ComputeCpuMax.java:26)",
                           "input": block("chapter2/README.md", 55, 57), "output": block("chapter2/README.md", 63, 65)}
    g["G4_avg"] = {"cite": "chapter2/README.md:152-164 (ComputeCpuAvg.java:31-58)",
                   "input": block("chapter2/README.md", 154, 157), "output": block("chapter2/README.md", 162, 163)}
    g["G5_median"] = {"cite": "chapter2/README.md:236-248 (ComputeCpuMiddle.java:36-47)",
                      "input": block("chapter2/README.md", 238, 241), "output": block("chapter2/README.md", 246, 247)}
    g["G6_sum"] = {"cite": "chapter3/README.md:71-81 (BandwidthMonitor.java:37-39)",
                   "input": block("chapter3/README.md", 72, 75), "output": ["(www.163.com,11200)"]}
    g["G7_event_time"] = {"cite": "chapter3/README.md:284-297 (BandwidthMonitorWithEventTime.java:30-55)",
                          "input": block("chapter3/README.md", 285, 289), "output": block("chapter3/README.md", 295, 296)}
    # sanity: the blocks are where SURVEY.md says they are
    assert g["G1_filter"]["output"] == ["2> (10.8.22.1,cpu2,99.2)"], g["G1_filter"]
    assert g["G3_rolling_max"]["output"][2] == "3> (10.8.22.1,cpu0,99.9)", g["G3_rolling_max"]
    assert g["G4_avg"]["output"] == ["3> 86.26666666666667", "3> 20.2"], g["G4_avg"]
    assert g["G5_median"]["output"] == ["3> 80.5", "3> 20.2"], g["G5_median"]
    assert len(g["G6_sum"]["input"]) == 4 and g["G6_sum"]["input"][3].endswith("1000"), g["G6_sum"]
    assert g["G7_event_time"]["output"][1] == "2> (www.163.com,0.0012969970703125)", g["G7_event_time"]
    with open(OUT, "w", encoding="utf-8") as f:
        json.dump(g, f, indent=1, ensure_ascii=False)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
