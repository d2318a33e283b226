# Each rank appends to one shared file from inside fluxmpi_println; rank order must hold.
import sys
import fluxmpi_b200 as fm

fm.Init()


class Appender:
    def __init__(self, path):
        self.path = path

    def write(self, s):
        if "hello" in s:
            with open(self.path, "a") as f:
                f.write(f"{fm.local_rank()} {s.strip()}\n")

    def flush(self):
        pass


for i in range(3):
    fm.fluxmpi_println("hello ", i, file=Appender(sys.argv[1]))
fm.Finalize()
