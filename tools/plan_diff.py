#!/usr/bin/env python3
"""Compare the services sections of two plan files (`m2k.plan`): one written here (move2kube_b200.planio.WritePlan)
and one written by `move2kube plan` on a machine that has Go (and, for CNB options, Docker).

    python tools/plan_diff.py OURS.plan THEIRS.plan [--ignore-cnb] [--ignore-repoinfo]

Both files are read in wire form (paths relative to rootDir / m2kassets, as stored), so the two machines'
absolute locations do not matter.  Exit status 0 = same services, 1 = differences (listed), 2 = unreadable.
Reference format: types/plan/planutils.go:164-201, types/plan/plan.go:186-215.
"""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main(argv):
    from move2kube_b200 import planio
    args = [a for a in argv if not a.startswith("--")]
    if len(args) != 2:
        print(__doc__)
        return 2
    plans = []
    for path in args:
        try:
            with open(path) as f:
                plans.append(planio.loads(f.read(), path))
        except Exception as e:  # noqa: BLE001 -- report and stop
            print("cannot read %s: %s" % (path, e), file=sys.stderr)
            return 2
    for p in plans:
        for name in list(p.Spec.Inputs.Services):
            lst = p.Spec.Inputs.Services[name]
            if "--ignore-cnb" in argv:
                lst = [s for s in lst if s.ContainerBuildType != "CNB"]
            if "--ignore-repoinfo" in argv:
                for s in lst:
                    s.RepoInfo = type(s.RepoInfo)()
            if lst:
                p.Spec.Inputs.Services[name] = lst
            else:
                del p.Spec.Inputs.Services[name]
    d = planio.diff_services(plans[0], plans[1])
    for name, i, field, a, b in d:
        print("%s[%d].%s: ours=%r theirs=%r" % (name, i, field, a, b))
    print("%d difference(s) in %d + %d service entries" % (
        len(d), sum(len(v) for v in plans[0].Spec.Inputs.Services.values()), sum(len(v) for v in plans[1].Spec.Inputs.Services.values())))
    return 1 if d else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
